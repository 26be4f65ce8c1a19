#!/bin/bash
# round 2, 2-GPU lease: plan-ahead.  EP tests, then the N=2 EP bench with and without it
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_ep.py -m gpu -q -x --timeout 600 > gpurun_out/r2k_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/r2k_pytest.log
tail -4 gpurun_out/r2k_pytest.log
port=29600
for flag in "" "--no-ep-plan-ahead"; do
  port=$((port+1))
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $port \
      bench.py --gpus 2 --steps 20 --warmup 3 $flag > gpurun_out/r2k_n2$flag.json 2> gpurun_out/r2k_n2$flag.err
  echo "flag=$flag exit $? $(python -c "import json,sys; d=json.loads(open('gpurun_out/r2k_n2$flag.json').read().strip().splitlines()[-1]); print(d['roofline']['us_per_layer'], d['value'], d['e2e']['value'], d['ep_parity']['f32_transport_vs_single_gpu_block'])" 2>&1 | tail -1)"
done
