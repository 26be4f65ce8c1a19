#!/bin/bash
# round 2, 2-GPU lease: whole GPU suite (f1, x_preq, drop-in over the reference model, TP on 2 GPUs, EP at world 1/2),
# then the EP bench at N=2 with the per-stage timeline (dispatch/combine kernels reworked for bytes in flight)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q --timeout 900 > gpurun_out/r2c_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/r2c_pytest.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus 2 --steps 20 --warmup 3 --ep-timeline > gpurun_out/r2c_bench_n2.json 2> gpurun_out/r2c_bench_n2.err
echo "bench n2 exit $?" >> gpurun_out/r2c_bench_n2.err
tail -25 gpurun_out/r2c_pytest.log; tail -c 1500 gpurun_out/r2c_bench_n2.json; tail -3 gpurun_out/r2c_bench_n2.err
