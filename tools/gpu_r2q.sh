#!/bin/bash
# round 2, 1-GPU lease: fused down projection, single-accumulator form + unordered pair groups — parity, then the sweep
# over token counts (chain / fold / one accumulator), then the headline line
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_tri.py -m gpu -q --timeout 300 > gpurun_out/r2q_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/r2q_pytest.log
tail -25 gpurun_out/r2q_pytest.log
timeout 400 python tools/tri_sweep.py > gpurun_out/r2q_tri_sweep.txt 2> gpurun_out/r2q_tri_sweep.err
echo "sweep exit $?"; cat gpurun_out/r2q_tri_sweep.txt; tail -5 gpurun_out/r2q_tri_sweep.err
timeout 200 python tools/tri_sweep.py --pw 0.95 --ns 65536,16384,8192 > gpurun_out/r2q_tri_sweep_pw95.txt 2>> gpurun_out/r2q_tri_sweep.err
cat gpurun_out/r2q_tri_sweep_pw95.txt
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-workloads > gpurun_out/r2q_bench.json 2> gpurun_out/r2q_bench.err
echo "bench exit $?"; python - <<PY
import json
d=json.loads(open('gpurun_out/r2q_bench.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['e2e']['value'], d.get('stages'), d['roofline'], d.get('block'))
PY
