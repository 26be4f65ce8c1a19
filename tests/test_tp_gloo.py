"""Tensor parallelism of the block on 2 ranks over gloo (CPU): the host logic — ``load_full_weights`` slicing
of I_e (layers/token_routed_mlp.py:252-258), the shared term pre-scaled by 1/tp (the shared expert is replicated and the
partial outputs are SUMMED), ONE all-reduce per block instead of the reference's K (token_routed_mlp.py:152) — with the
block's device compute replaced by the CPU oracle (test infrastructure only; the product has no CPU path).  The GPU/NCCL
version of the same check lives in tests/test_gpu_tp.py."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import i64_test_helpers as Hh
from oracle import i64_oracle as O


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


class _OracleBlock:
    """Stands in for kernels.block.I64Block on CPU: same constructor / load / forward surface, oracle arithmetic."""
    instances = []

    def __init__(self, hidden, num_experts, expert_inter, shared_inter, vocab, top_k, route_weights, rs, ss, mode, device):
        self.cfg = O.MoEConfig(hidden, num_experts, expert_inter, vocab, top_k=top_k,
                               primary_weight=route_weights[0] if top_k > 1 else 1.0, shared_inter=shared_inter,
                               shared_output_scale=ss, routed_output_scale=rs)
        self.ss = ss
        _OracleBlock.instances.append(self)

    def load_bf16(self, gate, up, down, sg=None, su=None, sd=None):
        f = lambda t: None if t is None else t.detach().float().numpy()
        self.w = O.MoEWeights(f(gate), f(up), f(down), f(sg), f(su), f(sd))

    def set_table(self, table):
        self.table = table.numpy()

    def forward(self, x, ids, out_dtype=torch.float32, out=None, stages=0xFF):
        y = O.moe_block_fp32(x.float().numpy(), None if ids is None else ids.numpy(), self.table, self.w, self.cfg)
        return torch.from_numpy(y).to(out_dtype)


def _worker(rank, world, port, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from vllm_i64 import _C
        from vllm_i64.parallel import tensor_parallel as TP
        TP.init_distributed(tp_size=world, backend="gloo")
        import vllm_i64.layers.token_routed_mlp as trm
        trm.I64Block = _OracleBlock                 # CPU stand-in for the device block (test only)
        _C.require_cuda = lambda *a, **k: None
        calls = {"n": 0}
        real_all_reduce = trm.all_reduce

        def counting_all_reduce(t):
            calls["n"] += 1
            return real_all_reduce(t)
        trm.all_reduce = counting_all_reduce

        h, e, ie, s, v, k, n = 64, 4, 32, 48, 300, 2, 41
        cfg, w, table, ids, x = Hh.make_block(h, e, ie, s, v, k, n, seed=3)
        m = trm.TokenRoutedMLP(h, e * ie, e, v, shared_expert=True, shared_intermediate_size=s, top_k=k,
                               top_k_primary_weight=0.5, shared_output_scale=cfg.shared_output_scale,
                               routed_output_scale=cfg.routed_output_scale)
        assert m.tp_size == world and m.expert_inter == ie // world
        assert tuple(m.gate_proj_w.shape) == (e, h, ie // world)
        m.load_full_weights(torch.from_numpy(w.gate), torch.from_numpy(w.up), torch.from_numpy(w.down))
        lo, hi = rank * (ie // world), (rank + 1) * (ie // world)
        np.testing.assert_array_equal(m.gate_proj_w.detach().numpy(), w.gate[:, :, lo:hi])
        np.testing.assert_array_equal(m.down_proj_w.detach().numpy(), w.down[:, lo:hi, :])
        with torch.no_grad():
            m.shared_gate.weight.copy_(torch.from_numpy(w.shared_gate))
            m.shared_up.weight.copy_(torch.from_numpy(w.shared_up))
            m.shared_down.weight.copy_(torch.from_numpy(w.shared_down))
            m.topk_token_to_expert.copy_(torch.from_numpy(table))
        out = m(torch.from_numpy(x), torch.from_numpy(ids))
        blk = _OracleBlock.instances[-1]
        assert abs(blk.ss - cfg.shared_output_scale / world) < 1e-7, "shared term must be pre-scaled by 1/tp"
        assert calls["n"] == 1, f"one all-reduce per block, saw {calls['n']}"
        ref = O.moe_block_fp32(x, ids, table, w, cfg)
        mx, mean = O.rel_errors(out.numpy(), ref)
        ret[rank] = (mx, mean)
    finally:
        dist.destroy_process_group()


def test_tp2_block_matches_unsharded_oracle():
    world = 2
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), ret), nprocs=world, join=True)
    for r in range(world):
        mx, mean = ret[r]
        assert mx < 1e-5 and mean < 1e-5, (r, mx, mean)       # fp32 sum of two half-width partial products
