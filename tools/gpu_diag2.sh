#!/bin/bash
# 2-GPU box: is the slowdown the box, the second visible GPU, NCCL at world 2, or P2P mappings?
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
show() { python - "$1" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    ep=d.get("expert_parallel") or {}
    print(sys.argv[1], round(d["value"]/1e6,2), "Mtok/s", round(d["ms_per_step"],4), "ms/step; e2e", round(d["e2e"]["ms_per_step"],4), "EP", ep.get("value"), (ep.get("nccl_all_to_all_baseline") or {}).get("value"))
except Exception as e: print(sys.argv[1], "parse fail", e)
PY
}
A="--steps 100 --warmup 10 --no-stages --no-cpu-baseline"
nvidia-smi --query-gpu=index,clocks.sm,clocks.mem,power.limit,persistence_mode,mig.mode.current,ecc.mode.current --format=csv
timeout 200 python bench.py $A > gpurun_out/d2_plain.json 2> gpurun_out/d2_plain.err; show gpurun_out/d2_plain.json
CUDA_VISIBLE_DEVICES=1 timeout 200 python bench.py $A > gpurun_out/d2_dev1.json 2> gpurun_out/d2_dev1.err; show gpurun_out/d2_dev1.json
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 200 $TR --master-port 29541 bench.py --gpus 2 $A --no-ep > gpurun_out/d2_t2.json 2> gpurun_out/d2_t2.err; show gpurun_out/d2_t2.json
NCCL_P2P_DISABLE=1 timeout 200 $TR --master-port 29542 bench.py --gpus 2 $A --no-ep > gpurun_out/d2_t2nop2p.json 2> gpurun_out/d2_t2nop2p.err; show gpurun_out/d2_t2nop2p.json
timeout 300 $TR --master-port 29544 bench.py --gpus 2 $A > gpurun_out/d2_t2ep.json 2> gpurun_out/d2_t2ep.err; show gpurun_out/d2_t2ep.json
