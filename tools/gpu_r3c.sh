#!/bin/bash
# 1-GPU lease: pair sort fused into the plan launch — parity (all fused-down-projection tests), then A/B lines
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_tri.py tests/test_gpu_fp8.py tests/test_gpu_host_pipeline.py -m gpu -q --timeout 600 > gpurun_out/r3c_pytest.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/r3c_pytest.log
for f in 1 0; do
for wl in ep64k_bf16 prefill_bf16 prefill_w8a8 prefill_fp8; do
I64B_TRI_PLAN_FUSED=$f timeout 300 python bench.py --workload $wl --steps 20 --warmup 3 --no-cpu-baseline --no-workloads --no-stages > gpurun_out/r3c_bench_${wl}_f$f.json 2> gpurun_out/r3c_bench.err
echo "fused=$f $wl exit $? $(python -c "import json; d=json.loads(open('gpurun_out/r3c_bench_${wl}_f$f.json').read().strip().splitlines()[-1]); print(round(d['value']/1e6,2), d['ms_per_step'], d['gpu_launches_per_step'])")"
done; done
tail -3 gpurun_out/r3c_bench.err
