#!/bin/bash
# 1-GPU lease, closing run of the round: smoke(), whole GPU suite, default bench line (timed), FP8 launch list after the fixes
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r3f_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/r3f_smoke.log
timeout 1500 python -m pytest tests -m gpu -q --timeout 900 > gpurun_out/r3f_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/r3f_pytest.log
tail -4 gpurun_out/r3f_pytest.log
/usr/bin/time -f "default bench wall %e s" timeout 900 python bench.py > gpurun_out/r3f_bench_n1.json 2> gpurun_out/r3f_bench_n1.err; echo "default bench rc=$?"; tail -1 gpurun_out/r3f_bench_n1.err
/usr/bin/time -f "reference arm wall %e s" timeout 600 python bench.py --impl reference > gpurun_out/r3f_bench_reference.json 2> gpurun_out/r3f_bench_reference.err; echo "reference arm rc=$?"; tail -1 gpurun_out/r3f_bench_reference.err
TAG=r3f
for wl in prefill_fp8 ep64k_bf16; do
CMD="python bench.py --workload $wl --steps 2 --warmup 3 --no-cpu-baseline --no-stages --no-workloads"
$CMD > gpurun_out/${TAG}_plain_$wl.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'i64b|plan_|tri_|dispatch_|combine_|requant|quantize' -c 800 --csv --log-file gpurun_out/${TAG}_launches_$wl.csv $CMD > gpurun_out/ncu_l_$wl.log 2>&1; echo "launches $wl rc=$?"
done
python tools/summarize_ncu.py $TAG > gpurun_out/${TAG}_summarize.log 2>&1
mkdir -p gpurun_out/${TAG}_profiles && cp profiles/${TAG}_* gpurun_out/${TAG}_profiles/ 2>/dev/null
du -sh gpurun_out
