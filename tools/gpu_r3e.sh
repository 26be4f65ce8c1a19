#!/bin/bash
# 1-GPU lease, timing-only probe: the pair GEMMs with a quarter of their operand bytes removed (no B / no A / neither on odd
# k-blocks; results are garbage) — is GEMM1 bound by the SMs' operand ingest?  Stage timings per workload and flag.
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
for wl in prefill_w8a8 prefill_fp8 ep64k_bf16; do
for f in 0 128 256 384; do
I64B_DEBUG_FLAGS=$f timeout 300 python bench.py --workload $wl --steps 10 --warmup 3 --no-cpu-baseline --no-workloads > gpurun_out/r3e_${wl}_f$f.json 2> gpurun_out/r3e.err
echo "$wl flags=$f $(python -c "import json; d=json.loads(open('gpurun_out/r3e_${wl}_f$f.json').read().strip().splitlines()[-1]); print({k:(v['us_per_launch'] if isinstance(v,dict) else v) for k,v in d['stages'].items() if k in ('gemm1','gemm2')})")"
done; done 2>&1 | tee gpurun_out/r3e_operand_probe.txt
tail -2 gpurun_out/r3e.err
