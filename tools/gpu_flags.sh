#!/bin/bash
# stage timing of one workload under several I64B_DEBUG_FLAGS values (bring-up only; results are wrong on purpose)
mkdir -p gpurun_out; export PYTHONDONTWRITEBYTECODE=1
WL=$1; shift
for f in "$@"; do
  I64B_DEBUG_FLAGS=$f timeout 300 python bench.py --workload $WL --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/flags_$f.json 2>gpurun_out/flags.err
  python - <<PY
import json
d=json.load(open("gpurun_out/flags_$f.json")); s=d["stages"]
print("flags=$f step_us/layer=%.1f" % s["_step_us_per_layer"], " ".join("%s=%.1f"%(k,v["us_per_launch"]) for k,v in s.items() if isinstance(v,dict)))
PY
done
