#!/bin/bash
# round 2, 1-GPU lease, final evidence of this round: whole GPU suite, the default bench line (all workloads) and the reference
# arm, then ncu launch lists + --set full captures of the headline (bf16 64K tokens: pair GEMM1, single-accumulator fused down
# projection), config 3 (W8A8: int8 GEMM1, re-quantiser, int8 fused down projection), the FP8 block and the decode kernel.
# Summarise with: python tools/summarize_ncu.py r2y
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
timeout 1500 python -m pytest tests -m gpu -q --timeout 900 > gpurun_out/r2y_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/r2y_pytest.log
tail -6 gpurun_out/r2y_pytest.log
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/r2y_bench_n1.json 2> gpurun_out/r2y_bench_n1.err; echo "default bench rc=$?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2y_bench_reference.json 2> gpurun_out/r2y_bench_reference.err; echo "reference arm rc=$?"
TAG=r2y
for wl in ep64k_bf16 prefill_w8a8 prefill_fp8 decode_bf16; do
CMD="python bench.py --workload $wl --steps 2 --warmup 3 --no-cpu-baseline --no-stages --no-workloads"
$CMD > gpurun_out/${TAG}_plain_$wl.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'i64b|plan_|tri_|dispatch_|combine_|requant|quantize' -c 800 --csv --log-file gpurun_out/${TAG}_launches_$wl.csv $CMD > gpurun_out/ncu_l_$wl.log 2>&1; echo "launches $wl rc=$?"
done
CMD="python bench.py --workload ep64k_bf16 --steps 2 --warmup 3 --no-cpu-baseline --no-stages --no-workloads"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'i64b_gemm|plan_|tri_|dispatch_kernel' -s 20 -c 10 -f -o gpurun_out/${TAG}_full_ep64k_bf16 $CMD > gpurun_out/ncu_f0.log 2>&1; echo "full ep64k rc=$?"
CMD="python bench.py --workload prefill_w8a8 --steps 2 --warmup 3 --no-cpu-baseline --no-stages --no-workloads"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'i64b_gemm|requant|dispatch_kernel' -s 20 -c 10 -f -o gpurun_out/${TAG}_full_prefill_w8a8 $CMD > gpurun_out/ncu_f1.log 2>&1; echo "full w8a8 rc=$?"
CMD="python bench.py --workload prefill_fp8 --steps 2 --warmup 3 --no-cpu-baseline --no-stages --no-workloads"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'i64b_gemm|requant|dispatch_kernel' -s 20 -c 10 -f -o gpurun_out/${TAG}_full_prefill_fp8 $CMD > gpurun_out/ncu_f2.log 2>&1; echo "full fp8 rc=$?"
CMD="python bench.py --workload decode_bf16 --steps 2 --warmup 3 --no-cpu-baseline --no-stages --no-workloads"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'i64b_fused' -s 30 -c 3 -f -o gpurun_out/${TAG}_full_decode_bf16 $CMD > gpurun_out/ncu_f3.log 2>&1; echo "full decode rc=$?"
# gpurun brings back at most 64 MiB: summarise on the box (profiles/r2y_*, profiles/ncu_traffic.json), keep the summaries
ls -la gpurun_out/*.ncu-rep | awk '{print $5, $9}'
python tools/summarize_ncu.py $TAG > gpurun_out/${TAG}_summarize.log 2>&1; echo "summarize rc=$?"
mkdir -p gpurun_out/${TAG}_profiles && cp profiles/${TAG}_* profiles/ncu_traffic.json gpurun_out/${TAG}_profiles/ 2>/dev/null
rm -f gpurun_out/*.ncu-rep
du -sh gpurun_out
tail -c 600 gpurun_out/r2y_bench_n1.json
