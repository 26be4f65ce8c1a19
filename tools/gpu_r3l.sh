#!/bin/bash
# 1-GPU lease: streamed end-to-end steps, number of micro-batches per step (1 = whole step per slot)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
for wl in ep64k_bf16 prefill_w8a8 decode_bf16; do
for ch in 1 2 3; do
timeout 300 python bench.py --workload $wl --steps 20 --warmup 5 --no-cpu-baseline --no-workloads --no-stages --e2e-chunks $ch > gpurun_out/r3l_bench_${wl}_ch$ch.json 2> gpurun_out/r3l_bench.err
echo "$wl chunks=$ch exit $? $(python -c "import json; d=json.loads(open('gpurun_out/r3l_bench_${wl}_ch$ch.json').read().strip().splitlines()[-1]); print(round(d['value']/1e6,2), round(d['e2e']['value']/1e6,2), d['e2e']['ms_per_step'])")"
done; done 2>&1 | tee gpurun_out/r3l_e2e_stream_chunks.txt
tail -2 gpurun_out/r3l_bench.err
