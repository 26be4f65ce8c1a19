#!/bin/bash
# round 2, 1-GPU lease: whole GPU suite, the default bench line, and the ncu evidence of the 64K-token headline with the
# token-major fused down projection (launch list + --set full of its kernels).  Summarise with tools/summarize_ncu.py r2o.
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
timeout 1500 python -m pytest tests -m gpu -q --timeout 900 > gpurun_out/r2o_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/r2o_pytest.log
tail -8 gpurun_out/r2o_pytest.log
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/r2o_bench_n1.json 2> gpurun_out/r2o_bench_n1.err; echo "default bench rc=$?"
TAG=r2o
wl=ep64k_bf16
CMD="python bench.py --workload $wl --steps 2 --warmup 3 --no-cpu-baseline --no-stages --no-workloads"
$CMD > gpurun_out/${TAG}_plain_$wl.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'i64b|plan_|tri_|dispatch_|combine_|requant|quantize' -c 600 --csv --log-file gpurun_out/${TAG}_launches_$wl.csv $CMD > gpurun_out/ncu_l_$wl.log 2>&1; echo "launches $wl rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'i64b_gemm|plan_|tri_|dispatch_kernel' -s 20 -c 10 -f -o gpurun_out/${TAG}_full_ep64k_bf16 $CMD > gpurun_out/ncu_f0.log 2>&1; echo "full ep64k rc=$?"
tail -c 700 gpurun_out/r2o_bench_n1.json
