#!/bin/bash
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
I64B_DEBUG_TIMELINE=1 timeout 300 python tools/probe_decode.py > gpurun_out/tl_0.log 2>&1; echo "rc=$?"; tail -32 gpurun_out/tl_0.log
I64B_PREFETCH=1 I64B_DEBUG_TIMELINE=1 timeout 300 python tools/probe_decode.py > gpurun_out/tl_1.log 2>&1; echo "rc=$?"; tail -12 gpurun_out/tl_1.log
