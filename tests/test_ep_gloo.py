"""Expert-parallel exchange logic (vllm_i64.parallel.expert_parallel) on 2 and 4 CPU ranks over gloo, with the
expert compute injected from the oracle: the per-rank outputs must equal the single-rank block on the same tokens
and the virtual-rank restatement of the reference's ``_forward_ep`` (layers/token_routed_i64/layer.py:209-298)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import i64_test_helpers as Hh
from oracle import i64_oracle as O


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, case, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from vllm_i64.parallel.expert_parallel import ep_moe_forward, init_expert_parallel, local_expert_range
        ep = init_expert_parallel()
        h, e, ie, s, v, k, n_total = case
        cfg, w, table, ids, x = Hh.make_block(h, e, ie, s, v, k, n_total, seed=3)
        cuts = np.linspace(0, n_total, world + 1).astype(int)
        if rank == world - 1 and world > 1:
            pass
        lo, hi = cuts[rank], cuts[rank + 1]
        xl, il = x[lo:hi], ids[lo:hi]
        routes = O.route_table(il, table, hi - lo, e)
        first, local_e = local_expert_range(e, ep)
        gate_up = np.concatenate([w.gate, w.up], axis=2)

        def local_fn(rows, local_ids):
            y = O.fused_token_routed_fp32(rows.numpy(), gate_up[first:first + local_e], w.down[first:first + local_e],
                                          local_ids.numpy().astype(np.int64), local_e)
            return torch.from_numpy(y)

        def shared_fn(xx):
            return torch.from_numpy(O.shared_forward_fp32(xx.numpy(), w))

        def combine(y, shared):
            yk = y.numpy().reshape(k, hi - lo, h)
            return torch.from_numpy(O.combine_fp32([yk[r] for r in range(k)], None if shared is None else shared.numpy(),
                                                   cfg))

        out = ep_moe_forward(torch.from_numpy(xl), torch.from_numpy(routes), e, local_fn, combine,
                             shared_fn if s else None, ep)
        ref = O.moe_block_fp32(xl, il, table, w, cfg)
        ret[rank] = float(np.abs(out.numpy() - ref).max() / max(np.abs(ref).max(), 1e-30))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,case", [(2, (64, 4, 16, 32, 300, 2, 101)), (4, (64, 8, 16, 0, 300, 1, 64)),
                                        (2, (64, 4, 16, 32, 300, 2, 3))])
def test_ep_forward_matches_single_rank(world, case):
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), case, ret), nprocs=world, join=True)
    assert len(ret) == world
    for r in range(world):
        assert ret[r] < 1e-5, (r, ret[r])


def test_exchange_plan_is_stable_by_rank():
    """Single-process check of the send order (stable by destination rank, then assignment index) against the
    virtual-rank oracle's ordering."""
    rng = np.random.RandomState(0)
    routes = rng.randint(0, 8, size=(2, 50))
    dest = routes.reshape(-1) // 2
    order = np.argsort(dest, kind="stable")
    t = torch.argsort(torch.from_numpy(dest), stable=True).numpy()
    np.testing.assert_array_equal(order, t)
    assert (np.diff(dest[order]) >= 0).all()


def _worker_batches(rank, world, port, equal, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from vllm_i64.parallel.expert_parallel import check_equal_local_batches, init_expert_parallel
        ep = init_expert_parallel()
        try:
            check_equal_local_batches(90 if equal else 87 + rank, ep, torch.device("cpu"))
            ret[rank] = "ok"
        except ValueError as exc:
            ret[rank] = str(exc)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("equal", [True, False])
def test_equal_local_batches_contract(equal):
    """The peer-memory exchange needs the same token count on every rank; a mismatch must raise on EVERY rank (a rank
    that alone walked into the id all-gather would hang: that is what a 700-token test split over 8 ranks once did)."""
    world = 2
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker_batches, args=(world, _free_port(), equal, ret), nprocs=world, join=True)
    for r in range(world):
        if equal:
            assert ret[r] == "ok"
        else:
            assert "same number of tokens" in ret[r] and "min 87, max 88" in ret[r], ret[r]
