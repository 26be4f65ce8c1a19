#!/bin/bash
# CTA-pair bring-up: parity with pairs forced on every block GEMM, then the default policy, then benches.
mkdir -p gpurun_out; export PYTHONDONTWRITEBYTECODE=1
I64B_CTA2=1 I64B_GEMM1_NMMA=256 I64B_GEMM2_NMMA=256 timeout 600 python -m pytest -x -q -p no:cacheprovider --timeout 120 --timeout-method=thread tests/test_gpu_parity.py -k "block or module or dense or determin" > gpurun_out/pairs_forced.log 2>&1
echo "forced rc=$?"; tail -n 5 gpurun_out/pairs_forced.log
SKIP_TESTS=${SKIP_TESTS_DEFAULT:-} TAG=${TAG:-p1} bash tools/gpu_iter.sh prefill_bf16 prefill_w8a8 dense_prefill
