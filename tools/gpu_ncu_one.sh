#!/bin/bash
# ncu --set full capture of a few launches of one kernel:  tools/gpu_ncu_one.sh <workload> <kernel-regex> <tag> [launch-skip] [count]
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
WL=$1; KREGEX=$2; TAG=$3; SKIP=${4:-8}; CNT=${5:-2}
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"$KREGEX" --launch-skip $SKIP --launch-count $CNT \
  -o gpurun_out/${TAG}_full_${WL} -f python bench.py --workload $WL --steps 2 --warmup 3 --no-cpu-baseline --no-stages > gpurun_out/${TAG}_ncu_${WL}.log 2>&1
echo "ncu rc=$?"; tail -n 3 gpurun_out/${TAG}_ncu_${WL}.log
