"""Three launches of the fused residual-add + RMSNorm + INT8 quant kernel at prefill size (for an ncu capture)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "vllm-i64_b200"))
import torch
from vllm_i64.kernels.cuda import get_i64_cuda_ops
ops = get_i64_cuda_ops()
dev = torch.device("cuda", 0)
n, h = 16384, 1024
x = torch.randn(n, h, device=dev).to(torch.bfloat16)
r = torch.randn(n, h, device=dev).to(torch.bfloat16)
w = torch.rand(h, device=dev).to(torch.bfloat16)
for _ in range(3):
    ops.add_rmsnorm(x, r, w, 1e-6, quantize=True, want_normed=True)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20):
    ops.add_rmsnorm(x, r, w, 1e-6, quantize=True, want_normed=True)
e1.record(); e1.synchronize()
us = e0.elapsed_time(e1) * 50
byts = n * h * (2 + 2 + 2 + 2 + 1) + n * 4
print(f"add_rmsnorm+quant n={n} h={h}: {us:.1f} us/launch, {byts / us * 1e-3:.0f} GB/s algorithmic")
