#!/bin/bash
# 1-GPU lease: uneven micro-batches in HostStepPipeline — parity, then the headline's e2e for several cuts
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_host_pipeline.py -m gpu -q --timeout 300 > gpurun_out/r3b_pytest.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/r3b_pytest.log
for ch in 4 1,2,2,2,1 1,3,4,4,3,1 1,2,3,4,3,2,1 2,3,3,3,3,2; do
tag=$(echo $ch | tr ',' '_')
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-workloads --no-stages --e2e-chunks $ch > gpurun_out/r3b_bench_ch$tag.json 2> gpurun_out/r3b_bench.err
echo "chunks=$ch exit $? $(python -c "import json; d=json.loads(open('gpurun_out/r3b_bench_ch$tag.json').read().strip().splitlines()[-1]); print(d['value'], d['e2e']['value'], d['e2e']['ms_per_step'])")"
done
tail -3 gpurun_out/r3b_bench.err
