#!/bin/bash
# Round-1 baseline call: GPU parity tests, smoke, bench lines, ncu launch list + one full capture of the top kernel.
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
timeout 1200 python -m pytest -x -q -p no:cacheprovider --timeout 300 --timeout-method=thread -m gpu tests > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 5 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -n 2 gpurun_out/smoke.log
for f in 1 0; do
  I64B_FUSED=$f timeout 400 python bench.py --workload decode_bf16 --steps 200 --warmup 10 --no-cpu-baseline > gpurun_out/r1a_decode_fused$f.json 2>gpurun_out/r1a_decode_fused$f.err; echo "decode fused=$f rc=$?"
done
for wl in prefill_bf16 prefill_w8a8 decode_w8a8 dense_prefill dense_decode tr_moe_306_decode; do
  timeout 400 python bench.py --workload $wl --steps 30 --warmup 10 --no-cpu-baseline > gpurun_out/r1a_$wl.json 2>gpurun_out/r1a_$wl.err; echo "$wl rc=$?"
done
timeout 600 python bench.py > gpurun_out/r1a_default.json 2>gpurun_out/r1a_default.err; echo "default rc=$?"
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-stages"
$CMD > gpurun_out/plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r1a_launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1; echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:i64b_fused_layer_kernel -s 30 -c 3 -f -o gpurun_out/r1a_fused $CMD > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
