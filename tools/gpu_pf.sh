#!/bin/bash
# cross-layer L2 prefetch experiment: decode bench with I64B_PREFETCH = 0 / 1 / 2, parity, timeline
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
for m in 0 1 2; do
  I64B_PREFETCH=$m timeout 300 python bench.py --no-cpu-baseline --no-stages > gpurun_out/pf_$m.json 2> gpurun_out/pf_$m.err; echo "pf $m rc=$?"
  python - <<PY
import json
d=json.loads(open("gpurun_out/pf_$m.json").read().strip().splitlines()[-1])
print("mode $m", round(d["value"]/1e6,2), "Mtok/s", d["roofline"]["us_per_launch"], "us/layer e2e", round(d["e2e"]["value"]/1e6,2))
PY
done
for wl in dense_decode tr_moe_306_decode; do
for m in 0 1; do
  I64B_PREFETCH=$m timeout 300 python bench.py --workload $wl --no-cpu-baseline --no-stages > gpurun_out/pf_${wl}_$m.json 2> gpurun_out/pf_${wl}_$m.err; echo "pf $wl $m rc=$?"
  python - <<PY
import json
d=json.loads(open("gpurun_out/pf_${wl}_$m.json").read().strip().splitlines()[-1])
print("$wl mode $m", round(d["value"]/1e6,2), "Mtok/s", d["ms_per_step"]*1000/d["config"]["layers_per_step"], "us/layer")
PY
done; done
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "fused or decode" 2>&1 | tail -3
I64B_DEBUG_TIMELINE=1 timeout 300 python tools/probe_decode.py > gpurun_out/pf_timeline.log 2>&1; echo "timeline rc=$?"; tail -30 gpurun_out/pf_timeline.log
