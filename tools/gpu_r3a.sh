#!/bin/bash
# 1-GPU lease: FP8 re-quantiser after the fix (bit-exactness + line), W8A8 GEMM1 with direct Q14 stores (I64B_DEBUG_FLAGS=64)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_fp8.py -m gpu -q --timeout 600 > gpurun_out/r3a_pytest_fp8.log 2>&1; echo "fp8 pytest exit $?"; tail -3 gpurun_out/r3a_pytest_fp8.log
I64B_DEBUG_FLAGS=64 timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_tri.py -m gpu -q --timeout 600 -k "w8a8" > gpurun_out/r3a_pytest_w8a8_direct.log 2>&1; echo "w8a8 direct pytest exit $?"; tail -3 gpurun_out/r3a_pytest_w8a8_direct.log
for f in 0 64; do
I64B_DEBUG_FLAGS=$f timeout 300 python bench.py --workload prefill_w8a8 --steps 20 --warmup 3 --no-cpu-baseline --no-workloads > gpurun_out/r3a_bench_w8a8_f$f.json 2> gpurun_out/r3a_bench.err
echo "flags=$f exit $? $(python -c "import json; d=json.loads(open('gpurun_out/r3a_bench_w8a8_f$f.json').read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], {k:(v['us_per_launch'] if isinstance(v,dict) else v) for k,v in d['stages'].items()})")"
done
timeout 300 python bench.py --workload prefill_fp8 --steps 20 --warmup 3 --no-cpu-baseline --no-workloads > gpurun_out/r3a_bench_fp8.json 2>> gpurun_out/r3a_bench.err
echo "fp8 exit $? $(python -c "import json; d=json.loads(open('gpurun_out/r3a_bench_fp8.json').read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], {k:(v['us_per_launch'] if isinstance(v,dict) else v) for k,v in d['stages'].items()})")"
tail -3 gpurun_out/r3a_bench.err
