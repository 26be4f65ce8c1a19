#!/bin/bash
# One GPU iteration: parity tests, then short benches of the named workloads (each under its own timeout).
#   tools/gpu_iter.sh [workload ...]      default: prefill_bf16 prefill_w8a8 dense_prefill decode_bf16
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
TAG=${TAG:-it}
if [ -z "$SKIP_TESTS" ]; then
  timeout 1200 python -m pytest -x -q -p no:cacheprovider --timeout 240 --timeout-method=thread tests/test_gpu_parity.py ${PYTEST_ARGS} > gpurun_out/${TAG}_pytest.log 2>&1
  echo "pytest rc=$?"; tail -n 8 gpurun_out/${TAG}_pytest.log
fi
show() { python - <<PY
import json
try:
    d=json.load(open("$1")); s=d["stages"]
    inner = s.get("unfused_chain_for_comparison", s)
    print("$2 value=%.4g e2e=%.4g step_us/layer=%.2f | " % (d["value"], d["e2e"]["value"], s["_step_us_per_layer"]) + " ".join("%s=%.1f" % (k, v["us_per_launch"]) for k, v in inner.items() if isinstance(v, dict)) + " | roof %s %.3f" % (d["roofline"]["kernel"], d["roofline"]["frac"]))
except Exception as e:
    print("$2 FAILED", e); print(open("gpurun_out/${TAG}.err").read()[-1500:])
PY
}
WLS=${@:-prefill_bf16 prefill_w8a8 dense_prefill decode_bf16}
for wl in $WLS; do
  steps=30; [ "$wl" = decode_bf16 ] && steps=200
  timeout 300 python bench.py --workload $wl --steps $steps --warmup 10 --no-cpu-baseline > gpurun_out/${TAG}_$wl.json 2>gpurun_out/${TAG}.err; show gpurun_out/${TAG}_$wl.json $wl
done
