#!/bin/bash
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
export I64B_GEMM1_NMMA=128 I64B_GEMM2_NMMA=128
I64B_DEBUG_TIMELINE=1 timeout 200 python tools/probe_decode.py > gpurun_out/probe_tl.log 2>&1; tail -n 125 gpurun_out/probe_tl.log
for f in 0 1 2 3; do
  I64B_DEBUG_FLAGS=$f timeout 300 python bench.py --workload decode_bf16 --steps 100 --warmup 10 --no-cpu-baseline > gpurun_out/dbg_$f.json 2>gpurun_out/dbg.err
  python - <<PY
import json
d=json.load(open("gpurun_out/dbg_$f.json")); s=d["stages"]
print("flags=$f gemm1=%.2f gemm2=%.2f" % (s["gemm1"]["us_per_launch"], s["gemm2"]["us_per_launch"]))
PY
done
