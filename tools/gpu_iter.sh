#!/bin/bash
# quick iteration: decode parity (fused-kernel tests) + decode bench (+ optional graph timeline)
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -3
timeout 300 python bench.py --no-cpu-baseline --no-stages > gpurun_out/it_decode.json 2> gpurun_out/it_decode.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open("gpurun_out/it_decode.json").read().strip().splitlines()[-1])
print("decode", round(d["value"]/1e6,2), "Mtok/s", d["roofline"]["us_per_launch"], "us/layer frac", d["roofline"]["frac"], "e2e", round(d["e2e"]["value"]/1e6,2))
PY
for wl in dense_decode tr_moe_306_decode; do
  timeout 300 python bench.py --workload $wl --no-cpu-baseline --no-stages > gpurun_out/it_$wl.json 2> gpurun_out/it_$wl.err
  python - <<PY
import json
d=json.loads(open("gpurun_out/it_$wl.json").read().strip().splitlines()[-1])
print("$wl", round(d["value"]/1e6,2), "Mtok/s", round(d["ms_per_step"]*1000/d["config"]["layers_per_step"],2), "us/layer")
PY
done
if [ "$1" = "tl" ]; then I64B_DEBUG_TIMELINE=1 timeout 300 python tools/probe_graph_tl.py 2>&1 | grep -v eager | tail -34; fi
