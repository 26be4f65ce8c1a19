#!/bin/bash
# 1-GPU lease: streaming end-to-end steps (HostStepPipeline.run_stream) — parity, then e2e streamed vs isolated
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_host_pipeline.py -m gpu -q --timeout 300 > gpurun_out/r3k_pytest.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/r3k_pytest.log
for flag in "" "--e2e-isolated" "--e2e-chunks 4" "--e2e-chunks 2"; do
tag=$(echo "$flag" | tr -d ' -')
timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-workloads --no-stages $flag > gpurun_out/r3k_bench_$tag.json 2> gpurun_out/r3k_bench.err
echo "flag='$flag' exit $? $(python -c "import json; d=json.loads(open('gpurun_out/r3k_bench_$tag.json').read().strip().splitlines()[-1]); print(round(d['value']/1e6,2), round(d['e2e']['value']/1e6,2), d['e2e']['ms_per_step'], d['e2e']['copies_per_step'])")"
done
for wl in prefill_w8a8 decode_bf16; do
timeout 300 python bench.py --workload $wl --steps 20 --warmup 5 --no-cpu-baseline --no-workloads --no-stages > gpurun_out/r3k_bench_$wl.json 2>> gpurun_out/r3k_bench.err
echo "$wl exit $? $(python -c "import json; d=json.loads(open('gpurun_out/r3k_bench_$wl.json').read().strip().splitlines()[-1]); print(round(d['value']/1e6,2), round(d['e2e']['value']/1e6,2), d['e2e']['ms_per_step'])")"
done
tail -3 gpurun_out/r3k_bench.err
