"""GEMM CTA-limit sweep on one GPU (bring-up tool): block time against i64b_moe_set_cta_limit for a few block shapes.
    python tools/cta_sweep.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "vllm-i64_b200")):
    sys.path.insert(0, p)
import torch
from vllm_i64 import _C
from vllm_i64.kernels.block import I64Block
from tools.synth_data import synthetic_hash_table

dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(0)
u = lambda shape, f: (torch.rand(shape, device=dev, generator=g) * 2 - 1) / f ** 0.5
V = 32000
CASES = [("shared-only S=1280", 1024, 0, 0, 1280, 0, 32768), ("shared-only S=1280", 1024, 0, 0, 1280, 0, 8192),
         ("500m block", 1024, 8, 512, 1280, 2, 65536), ("500m block", 1024, 8, 512, 1280, 2, 16384),
         ("routed-only 4 experts", 1024, 4, 512, 0, 1, 65536), ("dense-306", 1024, 0, 0, 4096, 0, 16384)]
for name, H, E, I, S, K, n in CASES:
    b = I64Block(H, E, I, S, V if E else 1, K, [0.5, 0.5][:K] if K else [], 2.0, 1.0, _C.MODE_BF16, dev)
    b.load_bf16(*( (u((E, H, I), H), u((E, H, I), H), u((E, I, H), I)) if E else (None, None, None)),
                *((u((S, H), H), u((S, H), H), u((H, S), S)) if S else (None, None, None)))
    if E:
        b.set_table(torch.from_numpy(synthetic_hash_table(V, E, K, 0)))
    x = torch.randn(n, H, device=dev).to(torch.bfloat16)
    ids = torch.randint(0, V, (n,), device=dev) if E else None
    out = torch.empty(n, H, dtype=torch.float32, device=dev)
    res = []
    for lim in (0, 146, 144, 140, 136, 132, 128, 120):
        b.set_cta_limit(lim)
        for _ in range(3):
            b.forward(x, ids, torch.float32, out=out)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            b.forward(x, ids, torch.float32, out=out)
        e1.record()
        e1.synchronize()
        res.append(f"{lim or 148}:{e0.elapsed_time(e1) * 100:.1f}")
    print(f"{name:24s} n={n:6d}  us/forward by CTA limit  " + "  ".join(res), flush=True)
    del b
