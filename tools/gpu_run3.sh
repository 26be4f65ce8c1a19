#!/bin/bash
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
timeout 900 python -m pytest -x -q -p no:cacheprovider --timeout 180 --timeout-method=thread tests/test_gpu_parity.py > gpurun_out/pytest.log 2>&1; echo "pytest rc=$?"; tail -n 6 gpurun_out/pytest.log
show() { python - <<PY
import json
d=json.load(open("$1")); s=d["stages"]
print("$2 value=%.3g e2e=%.3g step_us/layer=%.2f sum=%.2f | " % (d["value"], d["e2e"]["value"], s["_step_us_per_layer"], s["_sum_us_per_layer"]) + " ".join("%s=%.2f" % (k, v["us_per_launch"]) for k, v in s.items() if isinstance(v, dict)))
PY
}
for pdl in 1 0; do
  I64B_PDL=$pdl timeout 300 python bench.py --workload decode_bf16 --steps 200 --warmup 10 --no-cpu-baseline > gpurun_out/r3_decode_pdl$pdl.json 2>gpurun_out/r3.err; show gpurun_out/r3_decode_pdl$pdl.json "decode pdl=$pdl"
done
for g1 in 64 128 256; do
  I64B_GEMM1_NMMA=$g1 I64B_GEMM2_NMMA=$g1 timeout 300 python bench.py --workload decode_bf16 --steps 200 --warmup 10 --no-cpu-baseline > gpurun_out/r3_decode_$g1.json 2>gpurun_out/r3.err; show gpurun_out/r3_decode_$g1.json "decode nmma=$g1"
done
for wl in prefill_bf16 prefill_w8a8 decode_w8a8 dense_prefill; do
  timeout 300 python bench.py --workload $wl --steps 30 --warmup 10 --no-cpu-baseline > gpurun_out/r3_$wl.json 2>gpurun_out/r3.err; show gpurun_out/r3_$wl.json $wl
done
