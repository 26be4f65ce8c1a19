#!/bin/bash
# round 2, 2-GPU lease: NVLink access-pattern probe (tools/nvlink_probe.cu) + the torch.ops GPU tests
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 300 tools/_build/nvlink_probe > gpurun_out/r2e_nvlink_probe.txt 2>&1
echo "probe exit $?" >> gpurun_out/r2e_nvlink_probe.txt
timeout 600 python -m pytest tests/test_torch_ops.py -m gpu -q --timeout 300 > gpurun_out/r2e_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/r2e_pytest.log
cat gpurun_out/r2e_nvlink_probe.txt; tail -5 gpurun_out/r2e_pytest.log
