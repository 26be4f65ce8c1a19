#!/bin/bash
# 4-GPU box: EP parity at world 4 (even split) + the unequal-batch error, then the N = 4 bench line (clean box: nothing
# may be left running from a timed-out test before a bench)
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
timeout 110 python -m pytest -x -q -p no:cacheprovider --timeout 100 --timeout-method=thread "tests/test_gpu_ep.py::test_ep_block_matches_oracle[4]" tests/test_gpu_ep.py::test_ep_unequal_local_batches_raise_instead_of_hanging > gpurun_out/ep4_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/ep4_pytest.log | cut -c1-200
nvidia-smi --query-compute-apps=pid,used_memory --format=csv,noheader | head -5
timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29714 bench.py --gpus 4 --steps 50 --warmup 5 --no-stages > gpurun_out/ep4_bench_n4.json 2> gpurun_out/ep4_bench_n4.err; echo "bench n4 rc=$?"
python - <<'PY'
import json
try:
    d=json.loads(open("gpurun_out/ep4_bench_n4.json").read().strip().splitlines()[-1])
    ep=d.get("expert_parallel") or {}
    print("N=4 dp", round(d["value"]/1e6,2), "Mtok/s | EP", ep.get("value") and round(ep["value"]/1e6,2), "Mtok/s ms", ep.get("ms_per_step"), "nccl", (ep.get("nccl_all_to_all_baseline") or {}).get("value"), ep.get("error"))
except Exception as e:
    print("parse fail", e)
PY
