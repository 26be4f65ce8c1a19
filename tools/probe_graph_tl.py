"""Bring-up probe: per-CTA timeline of the last two fused decode layers INSIDE a CUDA-graph replay (PDL overlap between
layers included).  Run with I64B_DEBUG_TIMELINE=1."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "vllm-i64_b200")):
    sys.path.insert(0, p)
import torch
import bench
from vllm_i64 import _C
wl = dict(bench.WORKLOADS[sys.argv[1] if len(sys.argv) > 1 else "decode_bf16"]); wl["L"] = 8
dev = torch.device("cuda", 0)
layers = bench.build_layers(wl, dev)
x = torch.randn(wl["N"], wl["H"], device=dev).to(torch.bfloat16)
ids = torch.randint(0, wl["V"], (wl["N"],), device=dev)
outs = [torch.empty(wl["N"], wl["H"], dtype=torch.bfloat16, device=dev) for _ in range(2)]
stream = torch.cuda.Stream(dev)
def run():
    for i, b in enumerate(layers):
        b.forward(x, ids if wl["E"] else None, torch.bfloat16, out=outs[i & 1])
ms = bench.time_graph(run, 50, 10, stream)
print("us/layer", ms * 1e3 / (50 * wl["L"]))
_C.lib().i64b_debug_timeline_dump()
