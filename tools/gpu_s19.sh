#!/bin/bash
# session re-entry: parity + bench of every workload + ncu evidence for profiles/
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/s19_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/s19_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/s19_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/s19_smoke.log
timeout 600 python bench.py > gpurun_out/s19_bench_default.json 2> gpurun_out/s19_bench_default.err; echo "bench rc=$?"; cat gpurun_out/s19_bench_default.json
for wl in prefill_w8a8 prefill_bf16 dense_prefill decode_w8a8; do
  timeout 600 python bench.py --workload $wl --no-cpu-baseline > gpurun_out/s19_bench_$wl.json 2> gpurun_out/s19_bench_$wl.err; echo "bench $wl rc=$?"; cat gpurun_out/s19_bench_$wl.json
done
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/s19_bench_ref.json 2> gpurun_out/s19_bench_ref.err; echo "ref rc=$?"; cat gpurun_out/s19_bench_ref.json
bash tools/gpu_profiles.sh r1s19
