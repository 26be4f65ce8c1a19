#!/bin/bash
# round 2, 1-GPU lease: engine wiring test (reference loader passes, CUDAGraphRunner, histogram), FP8 bench line
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_dropin_reference.py -m gpu -q --timeout 600 > gpurun_out/r2w_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/r2w_pytest.log
grep -v "^$" gpurun_out/r2w_pytest.log | tail -30
timeout 300 python bench.py --workload prefill_fp8 --steps 20 --warmup 3 --no-cpu-baseline --no-workloads > gpurun_out/r2w_bench_fp8.json 2> gpurun_out/r2w_bench.err
echo "bench exit $?"; tail -3 gpurun_out/r2w_bench.err; python - <<PY
import json
d=json.loads(open('gpurun_out/r2w_bench_fp8.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d.get('stages'), d['roofline'], d.get('block'), d.get('measured_int8_peak'))
PY
