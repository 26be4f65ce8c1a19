"""Expert-parallel block on GPUs (NCCL): world_size 1 always, 2 / 4 / 8 when the box has that many devices.  Every rank's output
must match the single-GPU block on the same tokens (same kernels, so the match is tight) and the oracle."""
import datetime
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import i64_test_helpers as Hh
from oracle import i64_oracle as O

pytestmark = pytest.mark.gpu
if not torch.cuda.is_available():
    pytest.skip("CUDA device required", allow_module_level=True)


# a collective that cannot complete aborts the worker after this long (NCCL watchdog) instead of leaving it on the GPU:
# orphaned workers of a timed-out test once time-sliced every GPU of the box under the benches that followed
_PG_TIMEOUT = datetime.timedelta(seconds=120)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, case, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dev = torch.device("cuda", rank)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev, timeout=_PG_TIMEOUT)
    try:
        from vllm_i64 import _C
        from vllm_i64.parallel.expert_parallel import EPTokenRoutedBlock, init_expert_parallel
        ep = init_expert_parallel()
        h, e, ie, s, v, k, n_total = case
        cfg, w, table, ids, x = Hh.make_block(h, e, ie, s, v, k, n_total, seed=8)
        cuts = np.linspace(0, n_total, world + 1).astype(int)
        lo, hi = cuts[rank], cuts[rank + 1]
        t = lambda a, dt=None: (torch.from_numpy(np.ascontiguousarray(a)).to(dev) if dt is None
                                else torch.from_numpy(np.ascontiguousarray(a)).to(dev).to(dt))
        blk = EPTokenRoutedBlock(h, e, ie, s, v, k, cfg.route_weights(), cfg.routed_output_scale,
                                 cfg.shared_output_scale, _C.MODE_BF16, dev, ep)
        blk.load_bf16(t(w.gate), t(w.up), t(w.down), *((t(w.shared_gate), t(w.shared_up), t(w.shared_down)) if s else ()))
        blk.set_table(torch.from_numpy(table))
        xl, il = t(x[lo:hi], torch.bfloat16), t(ids[lo:hi])
        out = blk.forward(xl, il, torch.float32)                        # NVLink peer-memory exchange (product path)
        out_again = blk.forward(xl, il, torch.float32)                  # barrier epochs / buffer reuse
        out_nccl = blk.forward(xl, il, torch.float32, exchange="nccl")  # reference-style all_to_all_single exchange
        torch.cuda.synchronize(dev)
        assert torch.equal(out, out_again), "peer path is not run-to-run identical"
        assert torch.equal(out, out_nccl), "peer-memory and NCCL exchanges disagree"
        ref = O.moe_block_fp32(x[lo:hi], ids[lo:hi], table, w, cfg, inter_bf16=True)
        mx, mean = O.rel_errors(out.cpu().numpy(), ref)
        # bf16 transport of the expert outputs (the reference's EP semantics): one more bf16 rounding per assignment
        blk16 = EPTokenRoutedBlock(h, e, ie, s, v, k, cfg.route_weights(), cfg.routed_output_scale,
                                   cfg.shared_output_scale, _C.MODE_BF16, dev, ep, transport_dtype=torch.bfloat16)
        blk16.load_bf16(t(w.gate), t(w.up), t(w.down), *((t(w.shared_gate), t(w.shared_up), t(w.shared_down)) if s else ()))
        blk16.set_table(torch.from_numpy(table))
        out16 = blk16.forward(xl, il, torch.float32)
        torch.cuda.synchronize(dev)
        mx16, mean16 = O.rel_errors(out16.cpu().numpy(), ref)
        ret[rank] = (mx, mean, mx16, mean16)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [1, 2, 4, 8])
def test_ep_block_matches_oracle(world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    if world == 8 and os.environ.get("I64B_TEST_EP8") != "1":
        # 720 tokens split evenly over 1, 2, 4 and 8 ranks (the peer path needs equal local batches: with the earlier 700
        # tokens world 8 got 87/88-token ranks and hung in the id all-gather; EPTokenRoutedBlock now raises instead).
        # World 8 has not been re-run since that fix (GPU budget), so it stays opt-in until it has been seen green.
        pytest.skip("world 8 is opt-in (I64B_TEST_EP8=1)")
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), (256, 8, 64, 128, 1000, 2, 720), ret), nprocs=world, join=True)
    for r in range(world):
        mx, mean, mx16, mean16 = ret[r]
        assert mx <= Hh.TOL_MAX and mean <= Hh.TOL_MEAN, (r, mx, mean)
        # bf16 transport: per-assignment outputs rounded to bf16 (as the reference's EP path does); stated bar 2.5e-3
        assert mx16 <= Hh.TOL_MAX and mean16 <= 2.5e-3, (r, mx16, mean16)


def _worker_unequal(rank, world, port, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dev = torch.device("cuda", rank)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev, timeout=_PG_TIMEOUT)
    try:
        from vllm_i64 import _C
        from vllm_i64.parallel.expert_parallel import EPTokenRoutedBlock, init_expert_parallel
        ep = init_expert_parallel()
        h, e, ie, s, v, k = 256, 8, 64, 0, 1000, 2
        cfg, w, table, ids, x = Hh.make_block(h, e, ie, s, v, k, 64, seed=3)
        t = lambda a, dt=None: (torch.from_numpy(np.ascontiguousarray(a)).to(dev) if dt is None
                                else torch.from_numpy(np.ascontiguousarray(a)).to(dev).to(dt))
        blk = EPTokenRoutedBlock(h, e, ie, s, v, k, cfg.route_weights(), cfg.routed_output_scale,
                                 cfg.shared_output_scale, _C.MODE_BF16, dev, ep)
        blk.load_bf16(t(w.gate), t(w.up), t(w.down))
        blk.set_table(torch.from_numpy(table))
        n = 32 + rank                                   # ranks disagree on the batch size
        try:
            blk.forward(t(x[:n], torch.bfloat16), t(ids[:n]), torch.float32)
            ret[rank] = "no error"
        except ValueError as exc:
            ret[rank] = str(exc)
    finally:
        dist.destroy_process_group()


def test_ep_unequal_local_batches_raise_instead_of_hanging():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker_unequal, args=(2, _free_port(), ret), nprocs=2, join=True)
    for r in range(2):
        assert "same number of tokens" in ret[r], ret[r]
