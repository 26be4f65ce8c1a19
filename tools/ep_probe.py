"""Phase timing of the expert-parallel forward (bring-up tool): torchrun --nproc-per-node G tools/ep_probe.py"""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "vllm-i64_b200")):
    sys.path.insert(0, p)
import torch
import torch.distributed as dist
from vllm_i64 import _C
from vllm_i64.parallel.expert_parallel import EPTokenRoutedBlock, init_expert_parallel
from tools.synth_data import synthetic_hash_table

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
torch.cuda.set_device(dev)
dist.init_process_group("nccl", device_id=dev)
ep = init_expert_parallel()
H, E, I, S, V, K = 1024, 8, 512, 1280, 32000, 2
n = 65536 // world
g = torch.Generator(device=dev).manual_seed(0)
u = lambda shape, f: (torch.rand(shape, device=dev, generator=g) * 2 - 1) / f ** 0.5
b = EPTokenRoutedBlock(H, E, I, S, V, K, [0.5, 0.5], 2.0, 1.0, _C.MODE_BF16, dev, ep)
b.load_bf16(u((E, H, I), H), u((E, H, I), H), u((E, I, H), I), u((S, H), H), u((S, H), H), u((H, S), S))
b.set_table(torch.from_numpy(synthetic_hash_table(V, E, K, 0)))
x = torch.randn(n, H, device=dev).to(torch.bfloat16)
ids = torch.randint(0, V, (n,), device=dev)
for _ in range(3):
    b.forward(x, ids)
lib = _C.lib()
ctx = C.byref(b._ctx)
st = _C.stream_ptr(dev)
names, evs = [], []
shared_buf = torch.empty(n, H, dtype=torch.float32, device=dev)
out = torch.empty(n, H, dtype=torch.bfloat16, device=dev)


def mark(name):
    e = torch.cuda.Event(enable_timing=True)
    e.record()
    names.append(name)
    evs.append(e)


for it in range(3):
    names.clear(); evs.clear()
    torch.cuda.synchronize(); dist.barrier()
    mark("start")
    if world > 1:
        dist.all_gather_into_tensor(b._ids_all, ids)
    mark("allgather")
    _C.check(lib.i64b_ep_dispatch(ctx, x.data_ptr(), b._ids_all.data_ptr(), b._table.data_ptr(), I, b._ws_plan.data_ptr(), b._ws_plan.numel(), st))
    mark("plan+dispatch")
    _C.check(lib.i64b_ep_signal(ctx, st))
    mark("signal")
    shared = shared_buf
    b.shared.forward(x, None, torch.float32, out=shared, stages=_C.STAGE_PLAN)
    mark("sh_plan")
    b.shared.forward(x, None, torch.float32, out=shared, stages=_C.STAGE_GEMM1)
    mark("sh_gemm1")
    b.shared.forward(x, None, torch.float32, out=shared, stages=_C.STAGE_GEMM2)
    mark("sh_gemm2")
    _C.check(lib.i64b_ep_wait(ctx, st))
    mark("wait1")
    _C.check(lib.i64b_moe_forward_dispatched(b.local._h, ctx, b._ws_plan.data_ptr(), b._ws_gemm.data_ptr(), b._ws_gemm.numel(), st))
    mark("gemms+return")
    _C.check(lib.i64b_ep_signal(ctx, st)); _C.check(lib.i64b_ep_wait(ctx, st))
    mark("barrier2")
    wts = (C.c_float * 4)(0.5, 0.5, 0, 0)
    _C.check(lib.i64b_ep_combine(ctx, b._ws_plan.data_ptr(), shared.data_ptr(), wts, 2.0, 1.0, out.data_ptr(), _C.dtype_code(out), st))
    mark("combine")
    torch.cuda.synchronize()
    if rank == 0 and it == 2:
        print(f"EP{world} n_local={n}: " + "  ".join(f"{names[i]}={evs[i-1].elapsed_time(evs[i])*1e3:.0f}us" for i in range(1, len(evs)))
              + f"  total={evs[0].elapsed_time(evs[-1])*1e3:.0f}us")
dist.barrier()
dist.destroy_process_group()
