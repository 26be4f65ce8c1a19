#!/bin/bash
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
for wl in decode_bf16 prefill_w8a8 prefill_bf16 decode_w8a8 dense_decode dense_prefill; do
  st=300; [ "${wl#prefill}" != "$wl" -o "$wl" = dense_prefill ] && st=30
  timeout 600 python bench.py --workload $wl --steps $st --warmup 10 --no-cpu-baseline > gpurun_out/bench_$wl.json 2> gpurun_out/bench_$wl.err
  echo "== $wl rc=$?"; tail -c 3000 gpurun_out/bench_$wl.json; tail -n 5 gpurun_out/bench_$wl.err
done
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-stages"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_decode.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:i64b_gemm_kernel -s 2 -c 4 -o gpurun_out/prof_gemm_decode -f $CMD > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
tail -n 5 gpurun_out/ncu_full.log
