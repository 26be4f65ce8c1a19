#!/bin/bash
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
for g1 in 64 128 256; do for g2 in 64 128 256; do
  I64B_GEMM1_NMMA=$g1 I64B_GEMM2_NMMA=$g2 timeout 300 python bench.py --workload decode_bf16 --steps 200 --warmup 10 --no-cpu-baseline > gpurun_out/sw_${g1}_${g2}.json 2>gpurun_out/sw.err
  python - <<PY
import json
d=json.load(open("gpurun_out/sw_${g1}_${g2}.json"))
s=d["stages"]
print("g1=$g1 g2=$g2 step_us/layer=%.2f gemm1=%.2f gemm2=%.2f plan=%.2f disp=%.2f comb=%.2f" % (s["_step_us_per_layer"], s["gemm1"]["us_per_launch"], s["gemm2"]["us_per_launch"], s["plan"]["us_per_launch"], s["dispatch"]["us_per_launch"], s["combine"]["us_per_launch"]))
PY
done; done
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-stages"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"gemm_kernel|plan_|dispatch_kernel|combine_kernel|requant" -c 300 --csv --log-file gpurun_out/launches_decode.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches rc=$?"
