#!/bin/bash
# round 2, 1-GPU lease: GEMM1 gathers its routed A rows from x (no dispatched copy) — parity, then A/B lines
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_tri.py -m gpu -q -x --timeout 300 -k "gather" > gpurun_out/r2t_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/r2t_pytest.log
tail -25 gpurun_out/r2t_pytest.log
for g in 1 0; do
for wl in ep64k_bf16 prefill_bf16; do
I64B_GATHER=$g timeout 300 python bench.py --workload $wl --steps 20 --warmup 3 --no-cpu-baseline --no-workloads --no-stages > gpurun_out/r2t_bench_${wl}_g$g.json 2> gpurun_out/r2t_bench.err
echo "gather=$g $wl exit $? $(python -c "import json; d=json.loads(open('gpurun_out/r2t_bench_${wl}_g$g.json').read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], d['e2e']['value'])")"
done; done
tail -3 gpurun_out/r2t_bench.err
