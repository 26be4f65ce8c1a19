#!/bin/bash
# second look at an 8-GPU box: topology + P2P bandwidth first, then a single-GPU sanity bench, EP parity at world 8, benches
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
nvidia-smi topo -m 2>&1 | head -14 > gpurun_out/ep8b_topo.log; head -11 gpurun_out/ep8b_topo.log | cut -c1-110
timeout 120 python tools/p2p_probe.py 2>&1 | tee gpurun_out/ep8b_p2p.log | tail -9
timeout 120 python bench.py --steps 50 --warmup 5 --no-stages --no-cpu-baseline > gpurun_out/ep8b_single.json 2> gpurun_out/ep8b_single.err
python - <<'PY'
import json
try:
    d=json.loads(open("gpurun_out/ep8b_single.json").read().strip().splitlines()[-1]); print("single GPU", round(d["value"]/1e6,2), "Mtok/s", d["roofline"]["us_per_launch"], "us/layer")
except Exception as e: print("single parse fail", e)
PY
I64B_TEST_EP8=1 timeout 300 python -m pytest -x -q -p no:cacheprovider --timeout 240 --timeout-method=thread "tests/test_gpu_ep.py::test_ep_block_matches_oracle[8]" > gpurun_out/ep8b_pytest.log 2>&1; echo "pytest ep8 rc=$?"; tail -n 3 gpurun_out/ep8b_pytest.log | cut -c1-200
for n in 8 4; do
  timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2961$n bench.py --gpus $n --steps 50 --warmup 5 --no-stages > gpurun_out/ep8b_bench_n$n.json 2> gpurun_out/ep8b_bench_n$n.err; echo "bench n$n rc=$?"
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/ep8b_bench_n$n.json").read().strip().splitlines()[-1])
    ep=d.get("expert_parallel") or {}
    print("N=$n dp", round(d["value"]/1e6,2), "Mtok/s | EP", ep.get("value") and round(ep["value"]/1e6,2), "Mtok/s ms", ep.get("ms_per_step"), "nccl", (ep.get("nccl_all_to_all_baseline") or {}).get("value"), ep.get("error"))
except Exception as e:
    print("parse fail", e)
PY
done
