"""Tensor parallelism of the block on 2 GPUs over NCCL: I_e sliced over the ranks, the shared term pre-scaled
by 1/tp, ONE fp32 all-reduce per block; every rank's output must match the unsharded oracle (bf16 and W8A8 modes)."""
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


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    dev = torch.device("cuda", rank)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev, timeout=datetime.timedelta(seconds=120))
    try:
        from vllm_i64.parallel import tensor_parallel as TP
        TP.init_distributed(tp_size=world, backend="nccl")
        from vllm_i64.layers.token_routed_mlp import TokenRoutedMLP
        from vllm_i64.layers.dense_mlp import DenseMLP
        h, e, ie, s, v, k, n = 256, 4, 128, 128, 1000, 2, 777
        cfg, w, table, ids, x = Hh.make_block(h, e, ie, s, v, k, n, seed=13)
        m = TokenRoutedMLP(h, e * ie, e, v, shared_expert=True, shared_intermediate_size=s, top_k=k, top_k_primary_weight=0.5,
                           shared_output_scale=cfg.shared_output_scale, routed_output_scale=cfg.routed_output_scale)
        m.load_full_weights(torch.from_numpy(w.gate), torch.from_numpy(w.up), torch.from_numpy(w.down))
        with torch.no_grad():
            m.shared_gate.weight.copy_(torch.from_numpy(w.shared_gate)); m.shared_up.weight.copy_(torch.from_numpy(w.shared_up))
            m.shared_down.weight.copy_(torch.from_numpy(w.shared_down)); m.topk_token_to_expert.copy_(torch.from_numpy(table))
        m = m.to(dev)
        xt, it = torch.from_numpy(x).to(dev), torch.from_numpy(ids).to(dev)
        out = m(xt, it)
        ref = O.moe_block_fp32(x, ids, table, w, cfg, inter_bf16=True)
        res = {"moe": O.rel_errors(out.cpu().numpy(), ref)}
        # routed experts alone + the all-reduce (expert_forward)
        r0 = O.route_table(ids, table, n, e)[0]
        ef = m.expert_forward(xt, torch.from_numpy(r0).to(dev))
        res["expert_forward"] = O.rel_errors(ef.cpu().numpy(), O.expert_forward_fp32(x, r0, w, e, inter_bf16=True))
        # dense FFN: gate/up column-parallel, down row-parallel
        torch.manual_seed(0)
        d = DenseMLP(h, 512)
        full = {kk: torch.randn(sh, generator=torch.Generator().manual_seed(i)) * 0.05
                for i, (kk, sh) in enumerate((("g", (512, h)), ("u", (512, h)), ("d", (h, 512))))}
        full = {kk: vv.to(torch.bfloat16).float() for kk, vv in full.items()}
        d.gate_proj.load_full_weight(full["g"]); d.up_proj.load_full_weight(full["u"]); d.down_proj.load_full_weight(full["d"])
        d = d.to(dev)
        yd = d(xt)
        res["dense"] = O.rel_errors(yd.cpu().numpy(), O.dense_swiglu_fp32(x, full["g"].numpy(), full["u"].numpy(), full["d"].numpy(), inter_bf16=True))
        ret[rank] = res
    finally:
        dist.destroy_process_group()


def test_tp2_on_gpus():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(2, _free_port(), ret), nprocs=2, join=True)
    for r in range(2):
        for name, (mx, mean) in ret[r].items():
            assert mx <= Hh.TOL_MAX and mean <= Hh.TOL_MEAN, (r, name, mx, mean)
