#!/bin/bash
# First-contact script for a fresh B200 box: each stage runs under its own timeout so one hang cannot eat the call.
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
run() { # name timeout cmd...
  local name=$1 t=$2; shift 2
  echo "=== $name" | tee -a gpurun_out/summary.txt
  timeout $t "$@" > gpurun_out/$name.log 2>&1
  echo "exit $?" | tee -a gpurun_out/summary.txt
  tail -n 12 gpurun_out/$name.log | tee -a gpurun_out/summary.txt
}
PT="python -m pytest -x -q -p no:cacheprovider --timeout 120 --timeout-method=thread tests/test_gpu_parity.py"
run t_route 300 $PT -k "route_modulo or dispatch_plan or scatter_gather or quantize"
run t_bf16gemm 300 $PT -k "bf16_gemm"
run t_int8gemm 300 $PT -k "int8_gemm or int8_linear or weight_only"
run t_block_bf16 400 $PT -k "block_bf16 or deterministic or linearity"
run t_block_w8a8 400 $PT -k "w8a8"
run t_module 400 $PT -k "module or dense or kernels_package or host_buffer or cuda_graph"
run smoke 200 python __graft_entry__.py smoke
