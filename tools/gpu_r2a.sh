#!/bin/bash
# round 2, first lease (2 GPUs): EP v2 parity at world 1/2, the rest of the GPU suite, bench N=1 and N=2
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r2a_gpus.txt 2>&1
timeout 1200 python -m pytest tests/test_gpu_ep.py -x -q --timeout 600 > gpurun_out/r2a_ep_pytest.log 2>&1
echo "ep pytest exit $?" >> gpurun_out/r2a_ep_pytest.log
timeout 900 python -m pytest tests -m gpu -x -q --timeout 600 --ignore tests/test_gpu_ep.py > gpurun_out/r2a_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/r2a_pytest.log
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/r2a_bench_n1.json 2> gpurun_out/r2a_bench_n1.err
echo "bench n1 exit $?" >> gpurun_out/r2a_bench_n1.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus 2 --steps 20 --warmup 3 --ep-nccl-baseline > gpurun_out/r2a_bench_n2.json 2> gpurun_out/r2a_bench_n2.err
echo "bench n2 exit $?" >> gpurun_out/r2a_bench_n2.err
tail -3 gpurun_out/r2a_ep_pytest.log gpurun_out/r2a_pytest.log; tail -c 600 gpurun_out/r2a_bench_n1.json; tail -c 1500 gpurun_out/r2a_bench_n2.json; tail -5 gpurun_out/r2a_bench_n2.err
