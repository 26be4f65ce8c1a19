"""Isolated timing of the expert-parallel exchange kernels (bring-up tool):

    python -m torch.distributed.run --nproc-per-node G --master-addr 127.0.0.1 tools/ep_probe.py

Config-5 geometry, 64K tokens in total.  After one full forward (so that every buffer holds real data), each kernel is timed
ALONE, 10 back-to-back launches between two events after a group barrier (nothing else runs on the GPU), for every exchange
partition size in I64B_EP_PROBE_CTAS (default "0,16,32"): plan + dispatch, combine.  Prints GB/s over NVLink (remote rows only)
and over HBM (all bytes the kernel touches locally).  I64B_EP_TMA=0 selects the load/store kernels.
"""
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
b.reserve(n, lanes=1)
x = torch.randn(n, H, device=dev).to(torch.bfloat16)
ids = torch.randint(0, V, (n,), device=dev)
for _ in range(2):
    b.forward(x, ids)
torch.cuda.synchronize()
lib = _C.lib()
lane = b._lanes[0]
ctx = C.byref(lane.ctx)
st = _C.stream_ptr(dev)
shared = lane.shared_out[:n]
out = torch.empty(n, H, dtype=torch.bfloat16, device=dev)
wts = (C.c_float * 4)(0.5, 0.5, 0, 0)
remote = n * K * (world - 1) / world * H * 2
REPS = 10


def timed(fn):
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(REPS):
        fn()
    e1.record()
    e1.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) * 1e3 / REPS], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t[0])


def dispatch():
    _C.check(lib.i64b_ep_dispatch(ctx, x.data_ptr(), ids.data_ptr(), b._table.data_ptr(), lane.ws_plan.data_ptr(),
                                  lane.ws_plan.numel(), st))


def combine():
    _C.check(lib.i64b_ep_combine(ctx, lane.ws_plan.data_ptr(), shared.data_ptr(), wts, 2.0, 1.0, out.data_ptr(),
                                 _C.dtype_code(out), st))


def gemms():
    _C.check(lib.i64b_moe_forward_dispatched(b.local._h, ctx, lane.ws_plan.data_ptr(), lane.ws_gemm.data_ptr(),
                                             lane.ws_gemm.numel(), st))


def shared_gemms():
    b.shared.forward(x, None, torch.float32, out=shared)


for ctas in [int(v) for v in os.environ.get("I64B_EP_PROBE_CTAS", "0,16,32").split(",")]:
    lane.ctx.comm_ctas = ctas
    td = timed(dispatch)
    tc = timed(combine)
    if rank == 0:
        hbm_d = n * H * 2 * (1 + K / world)                     # x read + local rows written
        hbm_c = n * H * (2 * K / world + 4 + 2) + remote        # local y rows + shared fp32 + out (+ peers reading our y)
        print(f"EP{world} n_local={n} ctas={ctas or 'all'}: plan+dispatch {td:7.1f} us  ({remote / td * 1e-3:6.1f} GB/s NVLink out, "
              f"{hbm_d / td * 1e-3:7.1f} GB/s local)   combine {tc:7.1f} us ({remote / tc * 1e-3:6.1f} GB/s NVLink in, "
              f"{hbm_c / tc * 1e-3:7.1f} GB/s local)", flush=True)
b.local.set_cta_limit(0)
b.shared.set_cta_limit(0)
tg, ts = timed(gemms), timed(shared_gemms)
b.local.set_cta_limit(132)
b.shared.set_cta_limit(132)
tg2, ts2 = timed(gemms), timed(shared_gemms)
if rank == 0:
    print(f"EP{world}: routed GEMMs {tg:.1f} us (132 CTAs: {tg2:.1f})   shared GEMMs {ts:.1f} us (132 CTAs: {ts2:.1f})", flush=True)
dist.barrier()
b.close()
dist.destroy_process_group()
