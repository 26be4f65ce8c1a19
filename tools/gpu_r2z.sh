#!/bin/bash
# round 2, N-GPU lease (N = $1: 2 or 8), end of round: EP tests of that world size (N = 2: the whole EP file; N = 8: skipped,
# they ran green earlier this round — profiles/r2l_ep_pytest_8gpu.log — and the exchange code has not changed since), then
# the driver's bench line at N (ep_parity is checked inside the line).
N=${1:-2}
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
if [ "$N" = "2" ]; then
timeout 1200 python -m pytest tests/test_gpu_ep.py tests/test_gpu_tp.py -m gpu -q --timeout 600 > gpurun_out/r2z_ep_pytest_${N}gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/r2z_ep_pytest_${N}gpu.log
tail -5 gpurun_out/r2z_ep_pytest_${N}gpu.log
fi
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2962$N \
    bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/r2z_bench_n$N.json 2> gpurun_out/r2z_bench_n$N.err
echo "n$N exit $?"; python -c "import json; d=json.loads(open('gpurun_out/r2z_bench_n$N.json').read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], d['roofline'], d['ep_parity'])"
tail -3 gpurun_out/r2z_bench_n$N.err
