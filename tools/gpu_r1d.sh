#!/bin/bash
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
timeout 900 python -m pytest -x -q -p no:cacheprovider --timeout 120 -m gpu tests/test_gpu_parity.py > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 4 gpurun_out/pytest_gpu.log
I64B_DEBUG_TIMELINE=1 timeout 120 python tools/probe_decode.py decode_bf16 > gpurun_out/timeline_new.log 2>&1; echo "timeline rc=$?"
for wl in decode_bf16 tr_moe_306_decode dense_decode; do
timeout 300 python bench.py --workload $wl --steps 200 --warmup 10 --no-cpu-baseline --no-stages > gpurun_out/r1d_$wl.json 2>gpurun_out/r1d_$wl.err; echo "$wl rc=$?"
done
