#!/bin/bash
# fused decode kernel bring-up: parity tests, then decode bench with the fused kernel on / off
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
nvidia-smi -L
timeout 900 python -m pytest -x -q -p no:cacheprovider --timeout 180 --timeout-method=thread tests/test_gpu_parity.py > gpurun_out/pytest.log 2>&1; echo "pytest rc=$?"; tail -n 15 gpurun_out/pytest.log
show() { python - <<PY
import json
d=json.load(open("$1")); s=d["stages"]
print("$2 value=%.3g e2e=%.3g step_us/layer=%.2f roof=%s" % (d["value"], d["e2e"]["value"], s["_step_us_per_layer"], d["roofline"]))
PY
}
for f in 1 0; do
  I64B_FUSED=$f timeout 300 python bench.py --workload decode_bf16 --steps 200 --warmup 10 --no-cpu-baseline > gpurun_out/r4_decode_fused$f.json 2>gpurun_out/r4.err; echo "rc=$?"; tail -n 3 gpurun_out/r4.err; show gpurun_out/r4_decode_fused$f.json "decode fused=$f"
done
for wl in tr_moe_306_decode dense_decode; do
  timeout 300 python bench.py --workload $wl --steps 200 --warmup 10 --no-cpu-baseline > gpurun_out/r4_$wl.json 2>gpurun_out/r4.err; echo "rc=$?"; show gpurun_out/r4_$wl.json $wl
done
