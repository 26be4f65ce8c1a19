#!/bin/bash
# why is the decode step 2.5x slower under torchrun on the 8-GPU box?  separate torchrun / NCCL init / OMP threads
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
show() { python - "$1" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], round(d["value"]/1e6,2), "Mtok/s", round(d["ms_per_step"],4), "ms/step; e2e", round(d["e2e"]["ms_per_step"],4))
except Exception as e: print(sys.argv[1], "parse fail", e)
PY
}
A="--steps 100 --warmup 10 --no-stages --no-cpu-baseline"
timeout 200 python bench.py $A > gpurun_out/dg_plain.json 2> gpurun_out/dg_plain.err; show gpurun_out/dg_plain.json
OMP_NUM_THREADS=1 timeout 200 python bench.py $A > gpurun_out/dg_omp1.json 2> gpurun_out/dg_omp1.err; show gpurun_out/dg_omp1.json
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 1 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 1 $A > gpurun_out/dg_trun.json 2> gpurun_out/dg_trun.err; show gpurun_out/dg_trun.json
I64B_BENCH_FORCE_DIST=1 timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 1 --master-addr 127.0.0.1 --master-port 29532 bench.py --gpus 1 $A > gpurun_out/dg_nccl.json 2> gpurun_out/dg_nccl.err; show gpurun_out/dg_nccl.json; tail -3 gpurun_out/dg_nccl.err
nproc; cat /sys/fs/cgroup/cpu.max 2>/dev/null; env | grep -i -E "nccl|omp|cuda" | head
