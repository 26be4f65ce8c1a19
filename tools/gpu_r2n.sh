#!/bin/bash
# round 2, 1-GPU lease: token-major fused down projection — parity against the chain, then the headline with and without it
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_tri.py -m gpu -q -x --timeout 300 > gpurun_out/r2n_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/r2n_pytest.log
tail -15 gpurun_out/r2n_pytest.log
for tri in 1 0; do
I64B_TRI=$tri timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-workloads > gpurun_out/r2n_bench_tri$tri.json 2> gpurun_out/r2n_bench_tri$tri.err
echo "tri=$tri exit $?"; python - <<PY
import json
d=json.loads(open('gpurun_out/r2n_bench_tri$tri.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d.get('stages_us'), d['roofline'])
PY
done
