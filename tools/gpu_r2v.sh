#!/bin/bash
# round 2, 1-GPU lease: FP8 host mirror (core/fp8.py surface, module switches) + the fixed tests, then FP8 bench lines
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_fp8.py tests/test_gpu_parity.py::test_block_linearity_full_size tests/test_gpu_ep.py::test_ep_config5_shape -m gpu -q --timeout 300 > gpurun_out/r2v_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/r2v_pytest.log
grep -v "^$" gpurun_out/r2v_pytest.log | tail -30
