#!/bin/bash
# round 2, 8-GPU lease: exchange kernels alone (tools/ep_probe.py), then the N=8 EP bench over lanes x exchange-partition SMs
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
I64B_EP_PROBE_CTAS=0,16,32 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29580 \
    tools/ep_probe.py > gpurun_out/r2j_probe_n8.txt 2> gpurun_out/r2j_probe_n8.err
grep "^EP" gpurun_out/r2j_probe_n8.txt
port=29581
for cfg in "1 16" "1 32" "1 0" "2 16" "2 32"; do
  set -- $cfg; lanes=$1; comm=$2
  port=$((port+1))
  I64B_EP_COMM_SMS=$comm timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $port \
      bench.py --gpus 8 --steps 10 --warmup 3 --ep-lanes $lanes --ep-timeline > gpurun_out/r2j_n8_l${lanes}_c${comm}.json 2> gpurun_out/r2j_n8_l${lanes}_c${comm}.err
  echo "lanes=$lanes comm=$comm exit $? $(python -c "import json,sys; d=json.loads(open('gpurun_out/r2j_n8_l${lanes}_c${comm}.json').read().strip().splitlines()[-1]); print(d['roofline']['us_per_layer'], d['value'], d['ep_parity']['f32_transport_vs_single_gpu_block'], (d['ep_timeline'] or {}).get('max_over_ranks_us'))" 2>&1 | tail -1)"
done
