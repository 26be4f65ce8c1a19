#!/bin/bash
# round 2, 2-GPU lease: micro-batch lanes.  EP tests, then the N=2 EP bench over lanes {1,2,4} x {bulk-copy, load/store} kernels
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_ep.py -m gpu -q --timeout 600 > gpurun_out/r2f_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/r2f_pytest.log
tail -6 gpurun_out/r2f_pytest.log
port=29520
for tma in 1 0; do for lanes in 1 2 4; do
  port=$((port+1))
  I64B_EP_TMA=$tma timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $port \
      bench.py --gpus 2 --steps 10 --warmup 3 --ep-lanes $lanes > gpurun_out/r2f_n2_tma${tma}_l${lanes}.json 2> gpurun_out/r2f_n2_tma${tma}_l${lanes}.err
  echo "tma=$tma lanes=$lanes exit $? $(python -c "import json,sys; d=json.loads(open('gpurun_out/r2f_n2_tma${tma}_l${lanes}.json').read().strip().splitlines()[-1]); print(d['roofline']['us_per_layer'], d['value'], d['ep_parity']['f32_transport_vs_single_gpu_block'])" 2>&1 | tail -1)"
done; done
