#!/bin/bash
# 8-GPU box: EP parity at world 1/2/4/8, then the scaling bench at N = 2, 4, 8 (each line also carries the EP leg)
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
nvidia-smi -L | wc -l
timeout 500 python -m pytest -x -q -p no:cacheprovider --timeout 240 --timeout-method=thread tests/test_gpu_ep.py > gpurun_out/ep8_pytest.log 2>&1; echo "pytest ep rc=$?"; tail -n 6 gpurun_out/ep8_pytest.log
for n in 8 4 2; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 100 --warmup 10 --no-stages > gpurun_out/ep8_bench_n$n.json 2> gpurun_out/ep8_bench_n$n.err; echo "bench n$n rc=$?"
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/ep8_bench_n$n.json").read().strip().splitlines()[-1])
    ep=d.get("expert_parallel") or {}
    print("N=$n dp value", round(d["value"]/1e6,2), "Mtok/s | EP", ep.get("value") and round(ep["value"]/1e6,2), "Mtok/s ms", ep.get("ms_per_step"), "nccl", (ep.get("nccl_all_to_all_baseline") or {}).get("value"), ep.get("error"))
except Exception as e:
    print("parse fail", e)
PY
  tail -n 3 gpurun_out/ep8_bench_n$n.err
done
