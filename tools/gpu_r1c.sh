#!/bin/bash
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
for fl in 0 1 3; do
I64B_FUSED=0 I64B_DEBUG_FLAGS=$fl I64B_DEBUG_TIMELINE=1 timeout 300 python tools/probe_decode.py decode_bf16 > gpurun_out/tl_unfused_fl$fl.log 2>&1; echo "fl=$fl rc=$?"
done
