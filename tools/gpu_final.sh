#!/bin/bash
# end-of-session evidence: full GPU test suite, smoke, default bench (+ other workloads), ncu launch lists and full captures
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
TAG=${1:-r1s20}
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${TAG}_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/${TAG}_smoke.log
timeout 600 python bench.py > gpurun_out/${TAG}_bench_decode_bf16.json 2> gpurun_out/${TAG}_bench_decode_bf16.err; echo "bench rc=$?"; cut -c1-900 gpurun_out/${TAG}_bench_decode_bf16.json
for wl in dense_decode tr_moe_306_decode decode_w8a8; do
  timeout 600 python bench.py --workload $wl --no-cpu-baseline > gpurun_out/${TAG}_bench_$wl.json 2> gpurun_out/${TAG}_bench_$wl.err; echo "bench $wl rc=$?"; cut -c1-200 gpurun_out/${TAG}_bench_$wl.json
done
CMD="python bench.py --workload decode_bf16 --steps 2 --warmup 3 --no-cpu-baseline --no-stages"
$CMD > gpurun_out/${TAG}_plain_decode_bf16.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'i64b|plan_|dispatch_|combine_|requant' -c 600 --csv --log-file gpurun_out/${TAG}_launches_decode_bf16.csv $CMD > gpurun_out/ncu_l_decode.log 2>&1; echo "launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:i64b_fused_layer_kernel -s 40 -c 2 -f -o gpurun_out/${TAG}_full_decode_bf16 $CMD > gpurun_out/ncu_f1.log 2>&1; echo "full decode rc=$?"
