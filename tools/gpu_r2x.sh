#!/bin/bash
# round 2, 1-GPU lease: FP8 fused down projection — parity, bench line
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_fp8.py tests/test_gpu_tri.py -m gpu -q --timeout 600 > gpurun_out/r2x_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/r2x_pytest.log
grep -v "^$" gpurun_out/r2x_pytest.log | tail -20
timeout 300 python bench.py --workload prefill_fp8 --steps 20 --warmup 3 --no-cpu-baseline --no-workloads > gpurun_out/r2x_bench_fp8.json 2> gpurun_out/r2x_bench.err
echo "bench exit $?"; tail -3 gpurun_out/r2x_bench.err; python - <<PY
import json
d=json.loads(open('gpurun_out/r2x_bench_fp8.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], {k:(v['us_per_launch'] if isinstance(v,dict) else v) for k,v in d.get('stages').items()}, d['roofline'], d.get('block'))
PY
