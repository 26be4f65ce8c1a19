#!/bin/bash
# round 2, 8-GPU lease: EP parity at world 8 (and 4), bench N=4 / N=8 with the per-stage timeline
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r2b_gpus.txt 2>&1
timeout 1500 python -m pytest tests/test_gpu_ep.py -x -q --timeout 600 -k "8] or matches_oracle[4]" > gpurun_out/r2b_ep_pytest.log 2>&1
echo "ep pytest exit $?" >> gpurun_out/r2b_ep_pytest.log
for n in 8 4; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n \
    bench.py --gpus $n --steps 20 --warmup 3 --ep-timeline > gpurun_out/r2b_bench_n$n.json 2> gpurun_out/r2b_bench_n$n.err
echo "bench n$n exit $?" >> gpurun_out/r2b_bench_n$n.err
done
tail -3 gpurun_out/r2b_ep_pytest.log; tail -c 1200 gpurun_out/r2b_bench_n8.json; tail -3 gpurun_out/r2b_bench_n8.err
