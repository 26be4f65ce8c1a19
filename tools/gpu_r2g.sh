#!/bin/bash
# round 2, 2-GPU lease: SM partition for the exchange kernels.  EP tests, then the N=2 EP bench over lanes x comm SMs
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_ep.py -m gpu -q -x --timeout 600 > gpurun_out/r2g_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/r2g_pytest.log
tail -4 gpurun_out/r2g_pytest.log
port=29540
for cfg in "1 0" "1 16" "1 32" "1 48" "2 16" "2 32" "2 48"; do
  set -- $cfg; lanes=$1; comm=$2
  port=$((port+1))
  I64B_EP_COMM_SMS=$comm timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $port \
      bench.py --gpus 2 --steps 10 --warmup 3 --ep-lanes $lanes --ep-timeline > gpurun_out/r2g_n2_l${lanes}_c${comm}.json 2> gpurun_out/r2g_n2_l${lanes}_c${comm}.err
  echo "lanes=$lanes comm=$comm exit $? $(python -c "import json,sys; d=json.loads(open('gpurun_out/r2g_n2_l${lanes}_c${comm}.json').read().strip().splitlines()[-1]); print(d['roofline']['us_per_layer'], d['value'], d['ep_parity']['f32_transport_vs_single_gpu_block'], (d['ep_timeline'] or {}).get('max_over_ranks_us'))" 2>&1 | tail -1)"
done
