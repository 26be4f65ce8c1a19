#!/bin/bash
# 1-GPU lease: the driver's two commands as the driver runs them (default flags), with their wall time
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
t0=$SECONDS; timeout 900 python bench.py > gpurun_out/r3g_bench_n1.json 2> gpurun_out/r3g_bench_n1.err; echo "default bench rc=$? wall $((SECONDS-t0)) s"
t0=$SECONDS; timeout 600 python bench.py --impl reference > gpurun_out/r3g_bench_reference.json 2> gpurun_out/r3g_bench_reference.err; echo "reference arm rc=$? wall $((SECONDS-t0)) s"
tail -c 300 gpurun_out/r3g_bench_n1.json
