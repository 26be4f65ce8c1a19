#!/bin/bash
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
nvidia-smi -L
timeout 600 python -m pytest -x -q -p no:cacheprovider --timeout 300 --timeout-method=thread tests/test_gpu_ep.py > gpurun_out/pytest_ep.log 2>&1; echo "pytest ep rc=$?"; tail -n 8 gpurun_out/pytest_ep.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 100 --warmup 10 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo "bench n2 rc=$?"; tail -c 1500 gpurun_out/bench_n2.json; tail -n 5 gpurun_out/bench_n2.err
