#!/bin/bash
# ncu evidence for the bench command: launch list (per-launch device time) + one full capture of the dominant kernel
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'i64b|plan_|dispatch_|combine_|requant' -c 800 --csv --log-file gpurun_out/r1e_launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1; echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:i64b_fused_layer_kernel -s 40 -c 2 -f -o gpurun_out/r1e_fused $CMD > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
