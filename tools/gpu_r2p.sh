#!/bin/bash
# round 2, 1-GPU lease: HostStepPipeline — parity, then the headline's end-to-end step for 1 / 2 / 4 / 8 micro-batches
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_host_pipeline.py -m gpu -q -x --timeout 300 > gpurun_out/r2p_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/r2p_pytest.log
tail -6 gpurun_out/r2p_pytest.log
for ch in 1 2 4 8; do
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-workloads --no-stages --e2e-chunks $ch > gpurun_out/r2p_bench_ch$ch.json 2> gpurun_out/r2p_bench_ch$ch.err
echo "chunks=$ch exit $? $(python -c "import json; d=json.loads(open('gpurun_out/r2p_bench_ch$ch.json').read().strip().splitlines()[-1]); print(d['value'], d['e2e']['value'], d['e2e']['ms_per_step'])")"
done
