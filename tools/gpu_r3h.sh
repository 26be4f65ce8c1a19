#!/bin/bash
# 1-GPU lease: three-accumulator fused down projection with pipelined TMEM loads in its epilogues — parity, then lines
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_tri.py tests/test_gpu_fp8.py -m gpu -q --timeout 600 > gpurun_out/r3h_pytest.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/r3h_pytest.log
for wl in prefill_w8a8 prefill_fp8; do
timeout 300 python bench.py --workload $wl --steps 20 --warmup 3 --no-cpu-baseline --no-workloads > gpurun_out/r3h_bench_$wl.json 2> gpurun_out/r3h_bench.err
echo "$wl exit $? $(python -c "import json; d=json.loads(open('gpurun_out/r3h_bench_$wl.json').read().strip().splitlines()[-1]); print(round(d['value']/1e6,2), d['ms_per_step'], {k:(v['us_per_launch'] if isinstance(v,dict) else v) for k,v in d['stages'].items()})")"
done
timeout 200 python tools/tri_sweep.py --pw 0.95 --ns 65536,16384 > gpurun_out/r3h_tri_sweep_pw95.txt 2>> gpurun_out/r3h_bench.err; cat gpurun_out/r3h_tri_sweep_pw95.txt
tail -3 gpurun_out/r3h_bench.err
