#!/bin/bash
# 1-GPU lease: ncu evidence of the FINAL tree (after the FP8 re-quantiser fix and the fused plan launch): launch lists and
# --set full summaries for the headline, config 3 (W8A8), FP8 and decode.  Summarised on the box (gpurun returns <= 64 MiB).
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
TAG=r3j
for wl in ep64k_bf16 prefill_w8a8 prefill_fp8 prefill_bf16 decode_bf16; do
CMD="python bench.py --workload $wl --steps 2 --warmup 3 --no-cpu-baseline --no-stages --no-workloads"
$CMD > gpurun_out/${TAG}_plain_$wl.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'i64b|plan_|tri_|dispatch_|combine_|requant|quantize' -c 800 --csv --log-file gpurun_out/${TAG}_launches_$wl.csv $CMD > gpurun_out/ncu_l_$wl.log 2>&1; echo "launches $wl rc=$?"
done
CMD="python bench.py --workload ep64k_bf16 --steps 2 --warmup 3 --no-cpu-baseline --no-stages --no-workloads"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'i64b_gemm|plan_|dispatch_kernel' -s 16 -c 8 -f -o gpurun_out/${TAG}_full_ep64k_bf16 $CMD > gpurun_out/ncu_f0.log 2>&1; echo "full ep64k rc=$?"
CMD="python bench.py --workload prefill_w8a8 --steps 2 --warmup 3 --no-cpu-baseline --no-stages --no-workloads"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'i64b_gemm|requant|dispatch_kernel|plan_' -s 20 -c 10 -f -o gpurun_out/${TAG}_full_prefill_w8a8 $CMD > gpurun_out/ncu_f1.log 2>&1; echo "full w8a8 rc=$?"
CMD="python bench.py --workload prefill_fp8 --steps 2 --warmup 3 --no-cpu-baseline --no-stages --no-workloads"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'i64b_gemm|requant|dispatch_kernel|plan_' -s 20 -c 10 -f -o gpurun_out/${TAG}_full_prefill_fp8 $CMD > gpurun_out/ncu_f2.log 2>&1; echo "full fp8 rc=$?"
python tools/summarize_ncu.py $TAG > gpurun_out/${TAG}_summarize.log 2>&1; echo "summarize rc=$?"
mkdir -p gpurun_out/${TAG}_profiles && cp profiles/${TAG}_* profiles/ncu_traffic.json gpurun_out/${TAG}_profiles/ 2>/dev/null
rm -f gpurun_out/*.ncu-rep
du -sh gpurun_out
