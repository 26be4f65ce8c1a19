#!/bin/bash
# round 2, 2-GPU lease: bulk-copy (TMA) dispatch/combine.  EP + widen + torch.ops tests (no -x: every test reports), then
# the EP bench at N=2 with the per-stage timeline (I64B_EP_TMA=0 re-runs it on the load/store kernels for comparison).
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_ep.py tests/test_gpu_widen.py tests/test_torch_ops.py -m gpu -q --timeout 600 > gpurun_out/r2d_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/r2d_pytest.log
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus 2 --steps 20 --warmup 3 --ep-timeline > gpurun_out/r2d_bench_n2.json 2> gpurun_out/r2d_bench_n2.err
echo "bench n2 exit $?" >> gpurun_out/r2d_bench_n2.err
tail -25 gpurun_out/r2d_pytest.log; tail -c 1200 gpurun_out/r2d_bench_n2.json; tail -3 gpurun_out/r2d_bench_n2.err
