#!/bin/bash
# round 2, 1-GPU lease: W8A8 form of the fused down projection — parity (chain + oracle, bit-exact), sweep, config 3 line
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_tri.py -m gpu -q --timeout 300 > gpurun_out/r2r_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/r2r_pytest.log
tail -25 gpurun_out/r2r_pytest.log
timeout 400 python tools/tri_sweep.py --workload prefill_w8a8 --ns 65536,32768,16384,8192 > gpurun_out/r2r_tri_sweep_w8a8.txt 2> gpurun_out/r2r_tri_sweep.err
echo "sweep exit $?"; cat gpurun_out/r2r_tri_sweep_w8a8.txt; tail -5 gpurun_out/r2r_tri_sweep.err
timeout 300 python bench.py --workload prefill_w8a8 --steps 20 --warmup 3 --no-cpu-baseline --no-workloads > gpurun_out/r2r_bench_w8a8.json 2> gpurun_out/r2r_bench.err
echo "bench exit $?"; python - <<PY
import json
d=json.loads(open('gpurun_out/r2r_bench_w8a8.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d.get('stages'), d['roofline'], d.get('block'))
PY
