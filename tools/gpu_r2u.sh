#!/bin/bash
# round 2, 1-GPU lease: FP8 block — parity first, then the whole GPU suite (the GEMM kernel's operand type became a 3-way switch)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_fp8.py -m gpu -q -s --timeout 300 > gpurun_out/r2u_pytest_fp8.log 2>&1
echo "pytest exit $?" >> gpurun_out/r2u_pytest_fp8.log
grep -v "^$" gpurun_out/r2u_pytest_fp8.log | tail -40
timeout 900 python -m pytest tests -m gpu -q --timeout 300 --deselect tests/test_gpu_fp8.py > gpurun_out/r2u_pytest_all.log 2>&1
echo "pytest exit $?" >> gpurun_out/r2u_pytest_all.log
tail -8 gpurun_out/r2u_pytest_all.log
