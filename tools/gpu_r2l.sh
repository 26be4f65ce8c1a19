#!/bin/bash
# round 2, 8-GPU lease: world-8 EP tests, then the EP bench at N=8 (with the per-stage timeline) and N=4
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_ep.py -m gpu -q -k "8" --timeout 300 > gpurun_out/r2l_ep_pytest_8gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/r2l_ep_pytest_8gpu.log
tail -4 gpurun_out/r2l_ep_pytest_8gpu.log
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29611 \
    bench.py --gpus 8 --steps 20 --warmup 3 --ep-timeline > gpurun_out/r2l_bench_n8.json 2> gpurun_out/r2l_bench_n8.err
echo "n8 exit $?"; tail -c 1500 gpurun_out/r2l_bench_n8.json
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29612 \
    bench.py --gpus 8 --steps 20 --warmup 3 --no-ep-plan-ahead > gpurun_out/r2l_bench_n8_noplanahead.json 2> gpurun_out/r2l_bench_n8_noplanahead.err
echo "n8 no-plan-ahead exit $?"; python -c "import json; d=json.loads(open('gpurun_out/r2l_bench_n8_noplanahead.json').read().strip().splitlines()[-1]); print(d['roofline']['us_per_layer'], d['value'])"
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29613 \
    bench.py --gpus 4 --steps 20 --warmup 3 --ep-timeline > gpurun_out/r2l_bench_n4.json 2> gpurun_out/r2l_bench_n4.err
echo "n4 exit $?"; tail -c 1200 gpurun_out/r2l_bench_n4.json
