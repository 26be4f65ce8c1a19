#!/bin/bash
# round 2 ncu evidence (1 GPU): the default bench line, per-launch device times of three workloads, and --set full captures of
# their kernels.  Every ncu run follows a plain run of the same command that exited 0.  Summarise with tools/summarize_ncu.py r2.
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
TAG=r2
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/${TAG}_bench_n1.json 2> gpurun_out/${TAG}_bench_n1.err; echo "default bench rc=$?"
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${TAG}_bench_reference.json 2> gpurun_out/${TAG}_bench_reference.err; echo "reference arm rc=$?"
for wl in ep64k_bf16 decode_bf16 prefill_w8a8; do
  CMD="python bench.py --workload $wl --steps 2 --warmup 3 --no-cpu-baseline --no-stages --no-workloads"
  $CMD > gpurun_out/${TAG}_plain_$wl.log 2>&1 && \
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'i64b|plan_|dispatch_|combine_|requant|quantize' -c 600 --csv --log-file gpurun_out/${TAG}_launches_$wl.csv $CMD > gpurun_out/ncu_l_$wl.log 2>&1; echo "launches $wl rc=$?"
done
CMD="python bench.py --workload ep64k_bf16 --steps 2 --warmup 3 --no-cpu-baseline --no-stages --no-workloads"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'i64b_gemm_kernel|plan_|dispatch_kernel|combine_kernel' -s 28 -c 14 -f -o gpurun_out/${TAG}_full_ep64k_bf16 $CMD > gpurun_out/ncu_f0.log 2>&1; echo "full ep64k rc=$?"
CMD="python bench.py --workload decode_bf16 --steps 2 --warmup 3 --no-cpu-baseline --no-stages --no-workloads"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:i64b_fused_layer_kernel -s 40 -c 2 -f -o gpurun_out/${TAG}_full_decode_bf16 $CMD > gpurun_out/ncu_f1.log 2>&1; echo "full decode rc=$?"
CMD="python bench.py --workload prefill_w8a8 --steps 2 --warmup 3 --no-cpu-baseline --no-stages --no-workloads"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'i64b_gemm_kernel|requant|quantize|dispatch_kernel|combine_kernel' -s 32 -c 16 -f -o gpurun_out/${TAG}_full_prefill_w8a8 $CMD > gpurun_out/ncu_f2.log 2>&1; echo "full w8a8 rc=$?"
python tools/norm_probe.py > gpurun_out/${TAG}_norm_probe.txt 2>&1 && \
timeout 300 ncu --set full --clock-control none --import-source on -k regex:add_rmsnorm -s 3 -c 2 -f -o gpurun_out/${TAG}_full_add_rmsnorm python tools/norm_probe.py > gpurun_out/ncu_f3.log 2>&1; echo "full norm rc=$?"
cat gpurun_out/${TAG}_norm_probe.txt; ls -la gpurun_out/*.ncu-rep; tail -c 600 gpurun_out/${TAG}_bench_n1.json
