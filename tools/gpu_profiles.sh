#!/bin/bash
# ncu evidence for profiles/: per-launch device times of the bench command, and one --set full capture of the dominant
# kernel of the headline workload (decode) and of the INT8 prefill workload.  Each ncu run follows a plain run of the
# same command that exited 0.
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
TAG=${1:-r1}
for wl in decode_bf16 prefill_w8a8 prefill_bf16; do
  CMD="python bench.py --workload $wl --steps 2 --warmup 3 --no-cpu-baseline --no-stages"
  $CMD > gpurun_out/${TAG}_plain_$wl.log 2>&1 && \
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'i64b|plan_|dispatch_|combine_|requant' -c 600 --csv --log-file gpurun_out/${TAG}_launches_$wl.csv $CMD > gpurun_out/ncu_l_$wl.log 2>&1; echo "launches $wl rc=$?"
done
CMD="python bench.py --workload decode_bf16 --steps 2 --warmup 3 --no-cpu-baseline --no-stages"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:i64b_fused_layer_kernel -s 40 -c 2 -f -o gpurun_out/${TAG}_full_decode_bf16 $CMD > gpurun_out/ncu_f1.log 2>&1; echo "full decode rc=$?"
CMD="python bench.py --workload prefill_w8a8 --steps 2 --warmup 3 --no-cpu-baseline --no-stages"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:i64b_gemm_kernel -s 16 -c 2 -f -o gpurun_out/${TAG}_full_prefill_w8a8 $CMD > gpurun_out/ncu_f2.log 2>&1; echo "full w8a8 rc=$?"
CMD="python bench.py --workload prefill_bf16 --steps 2 --warmup 3 --no-cpu-baseline --no-stages"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:i64b_gemm_kernel -s 16 -c 2 -f -o gpurun_out/${TAG}_full_prefill_bf16 $CMD > gpurun_out/ncu_f3.log 2>&1; echo "full bf16 rc=$?"
