"""Bring-up probe: per-CTA timeline of the grouped GEMMs on the decode workload (env I64B_DEBUG_TIMELINE=1)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "vllm-i64_b200")):
    sys.path.insert(0, p)
import torch
import bench
wl = dict(bench.WORKLOADS[sys.argv[1] if len(sys.argv) > 1 else "decode_bf16"]); wl["L"] = 6
dev = torch.device("cuda", 0)
layers = bench.build_layers(wl, dev)
x = torch.randn(wl["N"], wl["H"], device=dev).to(torch.bfloat16)
ids = torch.randint(0, wl["V"], (wl["N"],), device=dev)
out = torch.empty(wl["N"], wl["H"], dtype=torch.bfloat16, device=dev)
for rep in range(2):
    for b in layers:
        b.forward(x, ids if wl["E"] else None, torch.bfloat16, out=out)
torch.cuda.synchronize()
print("done")
