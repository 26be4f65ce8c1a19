#!/usr/bin/env python
"""Fused down projection: chain (GEMM2 + combine) vs three-accumulator fold vs one accumulator, per token count.

    python tools/tri_sweep.py [--ns 65536,32768,...] [--steps 20] [--workload prefill_bf16]

Times L-layer CUDA-graph replays (L distinct weight sets, bench.py's builders) and prints us/layer per form, the
largest difference from the chain's output and which form the automatic selection takes.  Regenerates the table behind
I64B_TRI_MIN_ROWS_* (include/i64b.h) and profiles/r2q_tri_sweep.txt."""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

bench._product_paths()
import torch  # noqa: E402
from vllm_i64 import _C  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ns", default="65536,32768,16384,8192,4096")
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--workload", default="prefill_bf16")
    ap.add_argument("--pw", type=float, default=None, help="route weight override (0.95: no power of two -> fold only)")
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    wl = dict(bench.WORKLOADS[args.workload])
    if args.pw is not None:
        wl["pw"] = args.pw
    layers = bench.build_layers(wl, dev)
    stream = torch.cuda.Stream(dev)
    forms = [("chain", _C.DEBUG_NO_TRI), ("fold", _C.DEBUG_TRI | _C.DEBUG_TRI_FOLD), ("one_acc", _C.DEBUG_TRI), ("auto", 0)]
    print(f"# {wl['desc']}  L={wl['L']}  pw={wl['pw']} rs={wl['rs']} ss={wl['ss']}")
    print(f"{'n':>7} " + " ".join(f"{f:>10}" for f, _ in forms) + "   auto_form  max|d| fold/one_acc vs chain (of max|out|)")
    for n in [int(v) for v in args.ns.split(",")]:
        g = torch.Generator(device=dev).manual_seed(0)
        x = torch.randn(n, wl["H"], device=dev, generator=g).to(torch.bfloat16)
        ids = torch.randint(0, wl["V"], (n,), device=dev, generator=torch.Generator(device=dev).manual_seed(1234))
        outs = [torch.empty(n, wl["H"], device=dev, dtype=torch.bfloat16) for _ in layers]
        res, ref, diffs = [], None, []
        for name, flags in forms:
            for b in layers:
                b.set_debug(_C.DEBUG_NO_FUSED | flags)

            def run():
                for b, o in zip(layers, outs):
                    b.forward(x, ids, torch.bfloat16, out=o)
            try:
                ms = bench.time_graph(run, args.steps, 3, stream)
                res.append(ms / args.steps / len(layers) * 1e3)
                o0 = outs[0].float().clone()
                if name == "chain":
                    ref = o0
                elif name != "auto":
                    diffs.append(((o0 - ref).abs().max() / ref.abs().max()).item())
            except Exception as e:  # noqa: BLE001
                res.append(float("nan"))
                print(f"# n={n} {name}: {type(e).__name__}: {e}", file=sys.stderr)
        form = layers[0].tri_form(n)
        print(f"{n:>7} " + " ".join(f"{v:>10.1f}" for v in res) + f"   {form:>9}  " + " ".join(f"{d:.2e}" for d in diffs), flush=True)


if __name__ == "__main__":
    main()
