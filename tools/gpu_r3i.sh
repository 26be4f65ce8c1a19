#!/bin/bash
# 1-GPU lease: what the driver runs at round end, on the final tree: smoke(), pytest -m gpu, bench (its flags), reference arm
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r3i_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/r3i_smoke.log
timeout 1500 python -m pytest tests/ -x -q -m gpu > gpurun_out/r3i_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/r3i_pytest.log
t0=$SECONDS; timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r3i_bench_n1.json 2> gpurun_out/r3i_bench_n1.err; echo "bench rc=$? wall $((SECONDS-t0)) s"
t0=$SECONDS; timeout 600 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/r3i_bench_reference.json 2> gpurun_out/r3i_bench_reference.err; echo "reference arm rc=$? wall $((SECONDS-t0)) s"
python - <<PY
import json
d=json.loads(open('gpurun_out/r3i_bench_n1.json').read().strip().splitlines()[-1])
r=json.loads(open('gpurun_out/r3i_bench_reference.json').read().strip().splitlines()[-1])
print('value', round(d['value']/1e6,2), 'e2e', round(d['e2e']['value']/1e6,2), 'roofline', d['roofline']['frac'], 'block', d['block'], 'clocks', d['clocks'])
print('cpu_baseline', d['cpu_baseline'])
print('reference arm', r['value'], r['steps'], r['ms_per_step'], r['cpu_baseline']['sample'])
print({k: round(w['value']/1e6,2) for k,w in d['workloads'].items()})
PY
