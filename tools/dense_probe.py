"""Timing of the dense (shared-expert) block alone: python tools/dense_probe.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "vllm-i64_b200")):
    sys.path.insert(0, p)
import torch
from vllm_i64 import _C
from vllm_i64.kernels.block import I64Block
dev = torch.device("cuda", 0)
H = 1024
for S in (1280, 4096):
    blk = I64Block(H, 0, 0, S, 1, 0, [], 1.0, 1.0, _C.MODE_BF16, dev)
    g = torch.Generator(device=dev).manual_seed(0)
    u = lambda shape, f: (torch.rand(shape, device=dev, generator=g) * 2 - 1) / f ** 0.5
    blk.load_bf16(None, None, None, u((S, H), H), u((S, H), H), u((H, S), S))
    for n in (16384, 65536):
        x = torch.randn(n, H, device=dev).to(torch.bfloat16)
        for dt in (torch.float32, torch.bfloat16):
            out = torch.empty(n, H, dtype=dt, device=dev)
            for stages, name in ((_C.STAGE_ALL, "all"), (_C.STAGE_GEMM1, "gemm1"), (_C.STAGE_GEMM2, "gemm2")):
                for _ in range(3):
                    blk.forward(x, None, dt, out=out, stages=stages)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(10):
                    blk.forward(x, None, dt, out=out, stages=stages)
                e1.record(); e1.synchronize()
                print(f"S={S} n={n} out={dt} {name}: {e0.elapsed_time(e1) * 100:.0f} us")
