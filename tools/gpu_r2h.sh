#!/bin/bash
# round 2, 2-GPU lease: exchange kernels timed alone (tools/ep_probe.py), bulk-copy and load/store versions
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
for tma in 1 0; do
I64B_EP_TMA=$tma timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2957$tma \
    tools/ep_probe.py > gpurun_out/r2h_probe_tma$tma.txt 2> gpurun_out/r2h_probe_tma$tma.err
echo "tma=$tma exit $?"; grep "^EP" gpurun_out/r2h_probe_tma$tma.txt; tail -3 gpurun_out/r2h_probe_tma$tma.err
done
