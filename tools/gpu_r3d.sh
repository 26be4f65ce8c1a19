#!/bin/bash
# 1-GPU lease: sixteen epilogue warps in the W8A8 / FP8 GEMM1 pair kernels — parity (W8A8 bit-exact, FP8, EP world 1), lines
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_tri.py tests/test_gpu_fp8.py tests/test_gpu_widen.py tests/test_gpu_ep.py -m gpu -q --timeout 600 -k "w8a8 or W8A8 or fp8 or int8 or dense or ep_block" > gpurun_out/r3d_pytest.log 2>&1; echo "pytest exit $?"; tail -5 gpurun_out/r3d_pytest.log
for wl in prefill_w8a8 prefill_fp8; do
timeout 300 python bench.py --workload $wl --steps 20 --warmup 3 --no-cpu-baseline --no-workloads > gpurun_out/r3d_bench_$wl.json 2> gpurun_out/r3d_bench.err
echo "$wl exit $? $(python -c "import json; d=json.loads(open('gpurun_out/r3d_bench_$wl.json').read().strip().splitlines()[-1]); print(round(d['value']/1e6,2), d['ms_per_step'], {k:(v['us_per_launch'] if isinstance(v,dict) else v) for k,v in d['stages'].items()}, d['roofline']['frac'])")"
done
tail -3 gpurun_out/r3d_bench.err
