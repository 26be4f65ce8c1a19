#!/bin/bash
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
timeout 300 ./gpurun_out/stream_probe > gpurun_out/stream_probe.log 2>&1; echo "probe rc=$?"
timeout 600 python -m pytest -x -q -p no:cacheprovider --timeout 300 -m gpu tests/test_gpu_parity.py > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/pytest_gpu.log
for pf in 1 0; do
I64B_PREFETCH_L2=$pf I64B_DEBUG_TIMELINE=1 timeout 300 python tools/probe_decode.py decode_bf16 > gpurun_out/timeline_pf$pf.log 2>&1; echo "timeline pf=$pf rc=$?"
I64B_PREFETCH_L2=$pf timeout 400 python bench.py --workload decode_bf16 --steps 200 --warmup 10 --no-cpu-baseline --no-stages > gpurun_out/r1b_decode_pf$pf.json 2>gpurun_out/r1b_decode_pf$pf.err; echo "decode pf=$pf rc=$?"
done
