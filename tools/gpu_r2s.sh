#!/bin/bash
# round 2, 1-GPU lease: dispatch row copies on a second stream under the shared expert's GEMM1 — parity, then A/B lines
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_tri.py tests/test_gpu_host_pipeline.py -m gpu -q --timeout 300 > gpurun_out/r2s_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/r2s_pytest.log
tail -15 gpurun_out/r2s_pytest.log
for ov in 1 0; do
for wl in ep64k_bf16 prefill_bf16; do
I64B_OVERLAP=$ov timeout 300 python bench.py --workload $wl --steps 20 --warmup 3 --no-cpu-baseline --no-workloads --no-stages > gpurun_out/r2s_bench_${wl}_ov$ov.json 2> gpurun_out/r2s_bench.err
echo "overlap=$ov $wl exit $? $(python -c "import json; d=json.loads(open('gpurun_out/r2s_bench_${wl}_ov$ov.json').read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], d['e2e']['value'])")"
done; done
tail -3 gpurun_out/r2s_bench.err
