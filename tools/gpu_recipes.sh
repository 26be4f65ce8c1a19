#!/bin/bash
# The GPU-box recipes behind profiles/ — each function is the exact command sequence of one lease, named by the tag its
# outputs carry (gpurun_out/<tag>_* -> profiles/<tag>_*).  Run as:
#     gpurun [--gpus N] -- 'bash tools/gpu_recipes.sh <tag> [args]'
# tools/README.md maps tags to the files they produced.
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out

# round 2, 2-GPU lease: NVLink access-pattern probe (tools/nvlink_probe.cu) + the torch.ops GPU tests
r2e() {
    timeout 300 tools/_build/nvlink_probe > gpurun_out/r2e_nvlink_probe.txt 2>&1
    echo "probe exit $?" >> gpurun_out/r2e_nvlink_probe.txt
    timeout 600 python -m pytest tests/test_torch_ops.py -m gpu -q --timeout 300 > gpurun_out/r2e_pytest.log 2>&1
    echo "pytest exit $?" >> gpurun_out/r2e_pytest.log
    cat gpurun_out/r2e_nvlink_probe.txt; tail -5 gpurun_out/r2e_pytest.log
}

# round 2, 2-GPU lease: micro-batch lanes.  EP tests, then the N=2 EP bench over lanes {1,2,4} x {bulk-copy, load/store} kernels
r2f() {
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
}

# round 2, 2-GPU lease: SM partition for the exchange kernels.  EP tests, then the N=2 EP bench over lanes x comm SMs
r2g() {
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
}

# round 2, 2-GPU lease: exchange kernels timed alone (tools/ep_probe.py), bulk-copy and load/store versions
r2h() {
    for tma in 1 0; do
    I64B_EP_TMA=$tma timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2957$tma \
        tools/ep_probe.py > gpurun_out/r2h_probe_tma$tma.txt 2> gpurun_out/r2h_probe_tma$tma.err
    echo "tma=$tma exit $?"; grep "^EP" gpurun_out/r2h_probe_tma$tma.txt; tail -3 gpurun_out/r2h_probe_tma$tma.err
    done
}

# round 2, 8-GPU lease: exchange kernels alone (tools/ep_probe.py), then the N=8 EP bench over lanes x exchange-partition SMs
r2j() {
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
}

# round 2, 2-GPU lease: plan-ahead.  EP tests, then the N=2 EP bench with and without it
r2k() {
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
}

# round 2, 8-GPU lease: world-8 EP tests, then the EP bench at N=8 (with the per-stage timeline) and N=4
r2l() {
    timeout 900 python -m pytest tests/test_gpu_ep.py -m gpu -q -k "8" --timeout 300 > gpurun_out/r2l_ep_pytest_8gpu.log 2>&1
    echo "pytest exit $?" >> gpurun_out/r2l_ep_pytest_8gpu.log
    tail -4 gpurun_out/r2l_ep_pytest_8gpu.log
    timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29611 \
        bench.py --gpus 8 --steps 20 --warmup 3 --ep-timeline > gpurun_out/r2l_bench_n8.json 2> gpurun_out/r2l_bench_n8.err
    echo "n8 exit $?"; tail -c 1500 gpurun_out/r2l_bench_n8.json
    timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29612 \
        bench.py --gpus 8 --steps 20 --warmup 3 --no-ep-plan-ahead > gpurun_out/r2l_bench_n8_noplanahead.json 2> gpurun_out/r2l_bench_n8_noplanahead.err
    echo "n8 no-plan-ahead exit $?"; python -c "import json; d=json.loads(open('gpurun_out/r2l_bench_n8_noplanahead.json').read().strip().splitlines()[-1]); print(d['roofline']['us_per_layer'], d['value'])"
    timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29613 \
        bench.py --gpus 4 --steps 20 --warmup 3 --ep-timeline > gpurun_out/r2l_bench_n4.json 2> gpurun_out/r2l_bench_n4.err
    echo "n4 exit $?"; tail -c 1200 gpurun_out/r2l_bench_n4.json
}

# round 2 ncu evidence (1 GPU): the default bench line, per-launch device times of three workloads, and --set full captures of
# their kernels.  Every ncu run follows a plain run of the same command that exited 0.  Summarise with tools/summarize_ncu.py r2.
r2m() {
    export PYTHONDONTWRITEBYTECODE=1
    TAG=r2
    timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/${TAG}_bench_n1.json 2> gpurun_out/${TAG}_bench_n1.err; echo "default bench rc=$?"
    timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${TAG}_bench_reference.json 2> gpurun_out/${TAG}_bench_reference.err; echo "reference arm rc=$?"
    for wl in ep64k_bf16 decode_bf16 prefill_w8a8; do
      CMD="python bench.py --workload $wl --steps 2 --warmup 3 --no-cpu-baseline --no-stages --no-workloads"
      $CMD > gpurun_out/${TAG}_plain_$wl.log 2>&1 && \
      timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'i64b|plan_|dispatch_|combine_|requant|quantize' -c 600 --csv --log-file gpurun_out/${TAG}_launches_$wl.csv $CMD > gpurun_out/ncu_l_$wl.log 2>&1; echo "launches $wl rc=$?"
    done
    CMD="python bench.py --workload ep64k_bf16 --steps 2 --warmup 3 --no-cpu-baseline --no-stages --no-workloads"
    timeout 600 ncu --set full --clock-control none --import-source on -k regex:'i64b_gemm_kernel|plan_|dispatch_kernel|combine_kernel' -s 28 -c 14 -f -o gpurun_out/${TAG}_full_ep64k_bf16 $CMD > gpurun_out/ncu_f0.log 2>&1; echo "full ep64k rc=$?"
    CMD="python bench.py --workload decode_bf16 --steps 2 --warmup 3 --no-cpu-baseline --no-stages --no-workloads"
    timeout 600 ncu --set full --clock-control none --import-source on -k regex:i64b_fused_layer_kernel -s 40 -c 2 -f -o gpurun_out/${TAG}_full_decode_bf16 $CMD > gpurun_out/ncu_f1.log 2>&1; echo "full decode rc=$?"
    CMD="python bench.py --workload prefill_w8a8 --steps 2 --warmup 3 --no-cpu-baseline --no-stages --no-workloads"
    timeout 600 ncu --set full --clock-control none --import-source on -k regex:'i64b_gemm_kernel|requant|quantize|dispatch_kernel|combine_kernel' -s 32 -c 16 -f -o gpurun_out/${TAG}_full_prefill_w8a8 $CMD > gpurun_out/ncu_f2.log 2>&1; echo "full w8a8 rc=$?"
    python tools/norm_probe.py > gpurun_out/${TAG}_norm_probe.txt 2>&1 && \
    timeout 300 ncu --set full --clock-control none --import-source on -k regex:add_rmsnorm -s 3 -c 2 -f -o gpurun_out/${TAG}_full_add_rmsnorm python tools/norm_probe.py > gpurun_out/ncu_f3.log 2>&1; echo "full norm rc=$?"
    cat gpurun_out/${TAG}_norm_probe.txt; ls -la gpurun_out/*.ncu-rep; tail -c 600 gpurun_out/${TAG}_bench_n1.json
}

# round 2, 1-GPU lease: token-major fused down projection — parity against the chain, then the headline with and without it
r2n() {
    timeout 600 python -m pytest tests/test_gpu_tri.py -m gpu -q -x --timeout 300 > gpurun_out/r2n_pytest.log 2>&1
    echo "pytest exit $?" >> gpurun_out/r2n_pytest.log
    tail -15 gpurun_out/r2n_pytest.log
    for tri in 1 0; do
    I64B_TRI=$tri timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-workloads > gpurun_out/r2n_bench_tri$tri.json 2> gpurun_out/r2n_bench_tri$tri.err
    echo "tri=$tri exit $?"; python - <<PY
import json
d=json.loads(open('gpurun_out/r2n_bench_tri$tri.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d.get('stages_us'), d['roofline'])
PY
    done
}

# round 2, 1-GPU lease: whole GPU suite, the default bench line, and the ncu evidence of the 64K-token headline with the
# token-major fused down projection (launch list + --set full of its kernels).  Summarise with tools/summarize_ncu.py r2o.
r2o() {
    export PYTHONDONTWRITEBYTECODE=1
    timeout 1500 python -m pytest tests -m gpu -q --timeout 900 > gpurun_out/r2o_pytest.log 2>&1
    echo "pytest exit $?" >> gpurun_out/r2o_pytest.log
    tail -8 gpurun_out/r2o_pytest.log
    timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/r2o_bench_n1.json 2> gpurun_out/r2o_bench_n1.err; echo "default bench rc=$?"
    TAG=r2o
    wl=ep64k_bf16
    CMD="python bench.py --workload $wl --steps 2 --warmup 3 --no-cpu-baseline --no-stages --no-workloads"
    $CMD > gpurun_out/${TAG}_plain_$wl.log 2>&1 && \
    timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'i64b|plan_|tri_|dispatch_|combine_|requant|quantize' -c 600 --csv --log-file gpurun_out/${TAG}_launches_$wl.csv $CMD > gpurun_out/ncu_l_$wl.log 2>&1; echo "launches $wl rc=$?"
    timeout 600 ncu --set full --clock-control none --import-source on -k regex:'i64b_gemm|plan_|tri_|dispatch_kernel' -s 20 -c 10 -f -o gpurun_out/${TAG}_full_ep64k_bf16 $CMD > gpurun_out/ncu_f0.log 2>&1; echo "full ep64k rc=$?"
    tail -c 700 gpurun_out/r2o_bench_n1.json
}

# round 2, 1-GPU lease: HostStepPipeline — parity, then the headline's end-to-end step for 1 / 2 / 4 / 8 micro-batches
r2p() {
    timeout 600 python -m pytest tests/test_gpu_host_pipeline.py -m gpu -q -x --timeout 300 > gpurun_out/r2p_pytest.log 2>&1
    echo "pytest exit $?" >> gpurun_out/r2p_pytest.log
    tail -6 gpurun_out/r2p_pytest.log
    for ch in 1 2 4 8; do
    timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-workloads --no-stages --e2e-chunks $ch > gpurun_out/r2p_bench_ch$ch.json 2> gpurun_out/r2p_bench_ch$ch.err
    echo "chunks=$ch exit $? $(python -c "import json; d=json.loads(open('gpurun_out/r2p_bench_ch$ch.json').read().strip().splitlines()[-1]); print(d['value'], d['e2e']['value'], d['e2e']['ms_per_step'])")"
    done
}

# round 2, 1-GPU lease: fused down projection, single-accumulator form + unordered pair groups — parity, then the sweep
# over token counts (chain / fold / one accumulator), then the headline line
r2q() {
    timeout 600 python -m pytest tests/test_gpu_tri.py -m gpu -q --timeout 300 > gpurun_out/r2q_pytest.log 2>&1
    echo "pytest exit $?" >> gpurun_out/r2q_pytest.log
    tail -25 gpurun_out/r2q_pytest.log
    timeout 400 python tools/tri_sweep.py > gpurun_out/r2q_tri_sweep.txt 2> gpurun_out/r2q_tri_sweep.err
    echo "sweep exit $?"; cat gpurun_out/r2q_tri_sweep.txt; tail -5 gpurun_out/r2q_tri_sweep.err
    timeout 200 python tools/tri_sweep.py --pw 0.95 --ns 65536,16384,8192 > gpurun_out/r2q_tri_sweep_pw95.txt 2>> gpurun_out/r2q_tri_sweep.err
    cat gpurun_out/r2q_tri_sweep_pw95.txt
    timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-workloads > gpurun_out/r2q_bench.json 2> gpurun_out/r2q_bench.err
    echo "bench exit $?"; python - <<PY
import json
d=json.loads(open('gpurun_out/r2q_bench.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['e2e']['value'], d.get('stages'), d['roofline'], d.get('block'))
PY
}

# round 2, 1-GPU lease: W8A8 form of the fused down projection — parity (chain + oracle, bit-exact), sweep, config 3 line
r2r() {
    timeout 600 python -m pytest tests/test_gpu_tri.py -m gpu -q --timeout 300 > gpurun_out/r2r_pytest.log 2>&1
    echo "pytest exit $?" >> gpurun_out/r2r_pytest.log
    tail -25 gpurun_out/r2r_pytest.log
    timeout 400 python tools/tri_sweep.py --workload prefill_w8a8 --ns 65536,32768,16384,8192 > gpurun_out/r2r_tri_sweep_w8a8.txt 2> gpurun_out/r2r_tri_sweep.err
    echo "sweep exit $?"; cat gpurun_out/r2r_tri_sweep_w8a8.txt; tail -5 gpurun_out/r2r_tri_sweep.err
    timeout 300 python bench.py --workload prefill_w8a8 --steps 20 --warmup 3 --no-cpu-baseline --no-workloads > gpurun_out/r2r_bench_w8a8.json 2> gpurun_out/r2r_bench.err
    echo "bench exit $?"; python - <<PY
import json
d=json.loads(open('gpurun_out/r2r_bench_w8a8.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d.get('stages'), d['roofline'], d.get('block'))
PY
}

# round 2, 1-GPU lease: dispatch row copies on a second stream under the shared expert's GEMM1 — parity, then A/B lines
r2s() {
    timeout 600 python -m pytest tests/test_gpu_tri.py tests/test_gpu_host_pipeline.py -m gpu -q --timeout 300 > gpurun_out/r2s_pytest.log 2>&1
    echo "pytest exit $?" >> gpurun_out/r2s_pytest.log
    tail -15 gpurun_out/r2s_pytest.log
    for ov in 1 0; do
    for wl in ep64k_bf16 prefill_bf16; do
    I64B_OVERLAP=$ov timeout 300 python bench.py --workload $wl --steps 20 --warmup 3 --no-cpu-baseline --no-workloads --no-stages > gpurun_out/r2s_bench_${wl}_ov$ov.json 2> gpurun_out/r2s_bench.err
    echo "overlap=$ov $wl exit $? $(python -c "import json; d=json.loads(open('gpurun_out/r2s_bench_${wl}_ov$ov.json').read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], d['e2e']['value'])")"
    done; done
    tail -3 gpurun_out/r2s_bench.err
}

# round 2, 1-GPU lease: GEMM1 gathers its routed A rows from x (no dispatched copy) — parity, then A/B lines
r2t() {
    timeout 600 python -m pytest tests/test_gpu_tri.py -m gpu -q -x --timeout 300 -k "gather" > gpurun_out/r2t_pytest.log 2>&1
    echo "pytest exit $?" >> gpurun_out/r2t_pytest.log
    tail -25 gpurun_out/r2t_pytest.log
    for g in 1 0; do
    for wl in ep64k_bf16 prefill_bf16; do
    I64B_GATHER=$g timeout 300 python bench.py --workload $wl --steps 20 --warmup 3 --no-cpu-baseline --no-workloads --no-stages > gpurun_out/r2t_bench_${wl}_g$g.json 2> gpurun_out/r2t_bench.err
    echo "gather=$g $wl exit $? $(python -c "import json; d=json.loads(open('gpurun_out/r2t_bench_${wl}_g$g.json').read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], d['e2e']['value'])")"
    done; done
    tail -3 gpurun_out/r2t_bench.err
}

# round 2, 1-GPU lease: FP8 block — parity first, then the whole GPU suite (the GEMM kernel's operand type became a 3-way switch)
r2u() {
    timeout 600 python -m pytest tests/test_gpu_fp8.py -m gpu -q -s --timeout 300 > gpurun_out/r2u_pytest_fp8.log 2>&1
    echo "pytest exit $?" >> gpurun_out/r2u_pytest_fp8.log
    grep -v "^$" gpurun_out/r2u_pytest_fp8.log | tail -40
    timeout 900 python -m pytest tests -m gpu -q --timeout 300 --deselect tests/test_gpu_fp8.py > gpurun_out/r2u_pytest_all.log 2>&1
    echo "pytest exit $?" >> gpurun_out/r2u_pytest_all.log
    tail -8 gpurun_out/r2u_pytest_all.log
}

# round 2, 1-GPU lease: FP8 host mirror (core/fp8.py surface, module switches) + the fixed tests, then FP8 bench lines
r2v() {
    timeout 600 python -m pytest tests/test_gpu_fp8.py tests/test_gpu_parity.py::test_block_linearity_full_size tests/test_gpu_ep.py::test_ep_config5_shape -m gpu -q --timeout 300 > gpurun_out/r2v_pytest.log 2>&1
    echo "pytest exit $?" >> gpurun_out/r2v_pytest.log
    grep -v "^$" gpurun_out/r2v_pytest.log | tail -30
}

# round 2, 1-GPU lease: engine wiring test (reference loader passes, CUDAGraphRunner, histogram), FP8 bench line
r2w() {
    timeout 900 python -m pytest tests/test_dropin_reference.py -m gpu -q --timeout 600 > gpurun_out/r2w_pytest.log 2>&1
    echo "pytest exit $?" >> gpurun_out/r2w_pytest.log
    grep -v "^$" gpurun_out/r2w_pytest.log | tail -30
    timeout 300 python bench.py --workload prefill_fp8 --steps 20 --warmup 3 --no-cpu-baseline --no-workloads > gpurun_out/r2w_bench_fp8.json 2> gpurun_out/r2w_bench.err
    echo "bench exit $?"; tail -3 gpurun_out/r2w_bench.err; python - <<PY
import json
d=json.loads(open('gpurun_out/r2w_bench_fp8.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d.get('stages'), d['roofline'], d.get('block'), d.get('measured_int8_peak'))
PY
}

# round 2, 1-GPU lease: FP8 fused down projection — parity, bench line
r2x() {
    timeout 900 python -m pytest tests/test_gpu_fp8.py tests/test_gpu_tri.py -m gpu -q --timeout 600 > gpurun_out/r2x_pytest.log 2>&1
    echo "pytest exit $?" >> gpurun_out/r2x_pytest.log
    grep -v "^$" gpurun_out/r2x_pytest.log | tail -20
    timeout 300 python bench.py --workload prefill_fp8 --steps 20 --warmup 3 --no-cpu-baseline --no-workloads > gpurun_out/r2x_bench_fp8.json 2> gpurun_out/r2x_bench.err
    echo "bench exit $?"; tail -3 gpurun_out/r2x_bench.err; python - <<PY
import json
d=json.loads(open('gpurun_out/r2x_bench_fp8.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], {k:(v['us_per_launch'] if isinstance(v,dict) else v) for k,v in d.get('stages').items()}, d['roofline'], d.get('block'))
PY
}

# round 2, 1-GPU lease, final evidence of this round: whole GPU suite, the default bench line (all workloads) and the reference
# arm, then ncu launch lists + --set full captures of the headline (bf16 64K tokens: pair GEMM1, single-accumulator fused down
# projection), config 3 (W8A8: int8 GEMM1, re-quantiser, int8 fused down projection), the FP8 block and the decode kernel.
# Summarise with: python tools/summarize_ncu.py r2y
r2y() {
    export PYTHONDONTWRITEBYTECODE=1
    timeout 1500 python -m pytest tests -m gpu -q --timeout 900 > gpurun_out/r2y_pytest.log 2>&1
    echo "pytest exit $?" >> gpurun_out/r2y_pytest.log
    tail -6 gpurun_out/r2y_pytest.log
    timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/r2y_bench_n1.json 2> gpurun_out/r2y_bench_n1.err; echo "default bench rc=$?"
    timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2y_bench_reference.json 2> gpurun_out/r2y_bench_reference.err; echo "reference arm rc=$?"
    TAG=r2y
    for wl in ep64k_bf16 prefill_w8a8 prefill_fp8 decode_bf16; do
    CMD="python bench.py --workload $wl --steps 2 --warmup 3 --no-cpu-baseline --no-stages --no-workloads"
    $CMD > gpurun_out/${TAG}_plain_$wl.log 2>&1 && \
    timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'i64b|plan_|tri_|dispatch_|combine_|requant|quantize' -c 800 --csv --log-file gpurun_out/${TAG}_launches_$wl.csv $CMD > gpurun_out/ncu_l_$wl.log 2>&1; echo "launches $wl rc=$?"
    done
    CMD="python bench.py --workload ep64k_bf16 --steps 2 --warmup 3 --no-cpu-baseline --no-stages --no-workloads"
    timeout 600 ncu --set full --clock-control none --import-source on -k regex:'i64b_gemm|plan_|tri_|dispatch_kernel' -s 20 -c 10 -f -o gpurun_out/${TAG}_full_ep64k_bf16 $CMD > gpurun_out/ncu_f0.log 2>&1; echo "full ep64k rc=$?"
    CMD="python bench.py --workload prefill_w8a8 --steps 2 --warmup 3 --no-cpu-baseline --no-stages --no-workloads"
    timeout 600 ncu --set full --clock-control none --import-source on -k regex:'i64b_gemm|requant|dispatch_kernel' -s 20 -c 10 -f -o gpurun_out/${TAG}_full_prefill_w8a8 $CMD > gpurun_out/ncu_f1.log 2>&1; echo "full w8a8 rc=$?"
    CMD="python bench.py --workload prefill_fp8 --steps 2 --warmup 3 --no-cpu-baseline --no-stages --no-workloads"
    timeout 600 ncu --set full --clock-control none --import-source on -k regex:'i64b_gemm|requant|dispatch_kernel' -s 20 -c 10 -f -o gpurun_out/${TAG}_full_prefill_fp8 $CMD > gpurun_out/ncu_f2.log 2>&1; echo "full fp8 rc=$?"
    CMD="python bench.py --workload decode_bf16 --steps 2 --warmup 3 --no-cpu-baseline --no-stages --no-workloads"
    timeout 600 ncu --set full --clock-control none --import-source on -k regex:'i64b_fused' -s 30 -c 3 -f -o gpurun_out/${TAG}_full_decode_bf16 $CMD > gpurun_out/ncu_f3.log 2>&1; echo "full decode rc=$?"
    # gpurun brings back at most 64 MiB: summarise on the box (profiles/r2y_*, profiles/ncu_traffic.json), keep the summaries
    ls -la gpurun_out/*.ncu-rep | awk '{print $5, $9}'
    python tools/summarize_ncu.py $TAG > gpurun_out/${TAG}_summarize.log 2>&1; echo "summarize rc=$?"
    mkdir -p gpurun_out/${TAG}_profiles && cp profiles/${TAG}_* profiles/ncu_traffic.json gpurun_out/${TAG}_profiles/ 2>/dev/null
    rm -f gpurun_out/*.ncu-rep
    du -sh gpurun_out
    tail -c 600 gpurun_out/r2y_bench_n1.json
}

# round 2, N-GPU lease (N = $1: 2 or 8), end of round: EP tests of that world size (N = 2: the whole EP file; N = 8: skipped,
# they ran green earlier this round — profiles/r2l_ep_pytest_8gpu.log — and the exchange code has not changed since), then
# the driver's bench line at N (ep_parity is checked inside the line).
r2z() {
    N=${1:-2}
    if [ "$N" = "2" ]; then
    timeout 1200 python -m pytest tests/test_gpu_ep.py tests/test_gpu_tp.py -m gpu -q --timeout 600 > gpurun_out/r2z_ep_pytest_${N}gpu.log 2>&1
    echo "pytest exit $?" >> gpurun_out/r2z_ep_pytest_${N}gpu.log
    tail -5 gpurun_out/r2z_ep_pytest_${N}gpu.log
    fi
    timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2962$N \
        bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/r2z_bench_n$N.json 2> gpurun_out/r2z_bench_n$N.err
    echo "n$N exit $?"; python -c "import json; d=json.loads(open('gpurun_out/r2z_bench_n$N.json').read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], d['roofline'], d['ep_parity'])"
    tail -3 gpurun_out/r2z_bench_n$N.err
}

# 1-GPU lease: FP8 re-quantiser after the fix (bit-exactness + line), W8A8 GEMM1 with direct Q14 stores (I64B_DEBUG_FLAGS=64)
r3a() {
    timeout 900 python -m pytest tests/test_gpu_fp8.py -m gpu -q --timeout 600 > gpurun_out/r3a_pytest_fp8.log 2>&1; echo "fp8 pytest exit $?"; tail -3 gpurun_out/r3a_pytest_fp8.log
    I64B_DEBUG_FLAGS=64 timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_tri.py -m gpu -q --timeout 600 -k "w8a8" > gpurun_out/r3a_pytest_w8a8_direct.log 2>&1; echo "w8a8 direct pytest exit $?"; tail -3 gpurun_out/r3a_pytest_w8a8_direct.log
    for f in 0 64; do
    I64B_DEBUG_FLAGS=$f timeout 300 python bench.py --workload prefill_w8a8 --steps 20 --warmup 3 --no-cpu-baseline --no-workloads > gpurun_out/r3a_bench_w8a8_f$f.json 2> gpurun_out/r3a_bench.err
    echo "flags=$f exit $? $(python -c "import json; d=json.loads(open('gpurun_out/r3a_bench_w8a8_f$f.json').read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], {k:(v['us_per_launch'] if isinstance(v,dict) else v) for k,v in d['stages'].items()})")"
    done
    timeout 300 python bench.py --workload prefill_fp8 --steps 20 --warmup 3 --no-cpu-baseline --no-workloads > gpurun_out/r3a_bench_fp8.json 2>> gpurun_out/r3a_bench.err
    echo "fp8 exit $? $(python -c "import json; d=json.loads(open('gpurun_out/r3a_bench_fp8.json').read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], {k:(v['us_per_launch'] if isinstance(v,dict) else v) for k,v in d['stages'].items()})")"
    tail -3 gpurun_out/r3a_bench.err
}

# 1-GPU lease: uneven micro-batches in HostStepPipeline — parity, then the headline's e2e for several cuts
r3b() {
    timeout 600 python -m pytest tests/test_gpu_host_pipeline.py -m gpu -q --timeout 300 > gpurun_out/r3b_pytest.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/r3b_pytest.log
    for ch in 4 1,2,2,2,1 1,3,4,4,3,1 1,2,3,4,3,2,1 2,3,3,3,3,2; do
    tag=$(echo $ch | tr ',' '_')
    timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-workloads --no-stages --e2e-chunks $ch > gpurun_out/r3b_bench_ch$tag.json 2> gpurun_out/r3b_bench.err
    echo "chunks=$ch exit $? $(python -c "import json; d=json.loads(open('gpurun_out/r3b_bench_ch$tag.json').read().strip().splitlines()[-1]); print(d['value'], d['e2e']['value'], d['e2e']['ms_per_step'])")"
    done
    tail -3 gpurun_out/r3b_bench.err
}

# 1-GPU lease: pair sort fused into the plan launch — parity (all fused-down-projection tests), then A/B lines
r3c() {
    timeout 900 python -m pytest tests/test_gpu_tri.py tests/test_gpu_fp8.py tests/test_gpu_host_pipeline.py -m gpu -q --timeout 600 > gpurun_out/r3c_pytest.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/r3c_pytest.log
    for f in 1 0; do
    for wl in ep64k_bf16 prefill_bf16 prefill_w8a8 prefill_fp8; do
    I64B_TRI_PLAN_FUSED=$f timeout 300 python bench.py --workload $wl --steps 20 --warmup 3 --no-cpu-baseline --no-workloads --no-stages > gpurun_out/r3c_bench_${wl}_f$f.json 2> gpurun_out/r3c_bench.err
    echo "fused=$f $wl exit $? $(python -c "import json; d=json.loads(open('gpurun_out/r3c_bench_${wl}_f$f.json').read().strip().splitlines()[-1]); print(round(d['value']/1e6,2), d['ms_per_step'], d['gpu_launches_per_step'])")"
    done; done
    tail -3 gpurun_out/r3c_bench.err
}

# 1-GPU lease: sixteen epilogue warps in the W8A8 / FP8 GEMM1 pair kernels — parity (W8A8 bit-exact, FP8, EP world 1), lines
r3d() {
    timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_tri.py tests/test_gpu_fp8.py tests/test_gpu_widen.py tests/test_gpu_ep.py -m gpu -q --timeout 600 -k "w8a8 or W8A8 or fp8 or int8 or dense or ep_block" > gpurun_out/r3d_pytest.log 2>&1; echo "pytest exit $?"; tail -5 gpurun_out/r3d_pytest.log
    for wl in prefill_w8a8 prefill_fp8; do
    timeout 300 python bench.py --workload $wl --steps 20 --warmup 3 --no-cpu-baseline --no-workloads > gpurun_out/r3d_bench_$wl.json 2> gpurun_out/r3d_bench.err
    echo "$wl exit $? $(python -c "import json; d=json.loads(open('gpurun_out/r3d_bench_$wl.json').read().strip().splitlines()[-1]); print(round(d['value']/1e6,2), d['ms_per_step'], {k:(v['us_per_launch'] if isinstance(v,dict) else v) for k,v in d['stages'].items()}, d['roofline']['frac'])")"
    done
    tail -3 gpurun_out/r3d_bench.err
}

# 1-GPU lease, timing-only probe: the pair GEMMs with a quarter of their operand bytes removed (no B / no A / neither on odd
# k-blocks; results are garbage) — is GEMM1 bound by the SMs' operand ingest?  Stage timings per workload and flag.
r3e() {
    for wl in prefill_w8a8 prefill_fp8 ep64k_bf16; do
    for f in 0 128 256 384; do
    I64B_DEBUG_FLAGS=$f timeout 300 python bench.py --workload $wl --steps 10 --warmup 3 --no-cpu-baseline --no-workloads > gpurun_out/r3e_${wl}_f$f.json 2> gpurun_out/r3e.err
    echo "$wl flags=$f $(python -c "import json; d=json.loads(open('gpurun_out/r3e_${wl}_f$f.json').read().strip().splitlines()[-1]); print({k:(v['us_per_launch'] if isinstance(v,dict) else v) for k,v in d['stages'].items() if k in ('gemm1','gemm2')})")"
    done; done 2>&1 | tee gpurun_out/r3e_operand_probe.txt
    tail -2 gpurun_out/r3e.err
}

# 1-GPU lease, closing run of the round: smoke(), whole GPU suite, default bench line (timed), FP8 launch list after the fixes
r3f() {
    export PYTHONDONTWRITEBYTECODE=1
    timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r3f_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/r3f_smoke.log
    timeout 1500 python -m pytest tests -m gpu -q --timeout 900 > gpurun_out/r3f_pytest.log 2>&1
    echo "pytest exit $?" >> gpurun_out/r3f_pytest.log
    tail -4 gpurun_out/r3f_pytest.log
    /usr/bin/time -f "default bench wall %e s" timeout 900 python bench.py > gpurun_out/r3f_bench_n1.json 2> gpurun_out/r3f_bench_n1.err; echo "default bench rc=$?"; tail -1 gpurun_out/r3f_bench_n1.err
    /usr/bin/time -f "reference arm wall %e s" timeout 600 python bench.py --impl reference > gpurun_out/r3f_bench_reference.json 2> gpurun_out/r3f_bench_reference.err; echo "reference arm rc=$?"; tail -1 gpurun_out/r3f_bench_reference.err
    TAG=r3f
    for wl in prefill_fp8 ep64k_bf16; do
    CMD="python bench.py --workload $wl --steps 2 --warmup 3 --no-cpu-baseline --no-stages --no-workloads"
    $CMD > gpurun_out/${TAG}_plain_$wl.log 2>&1 && \
    timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'i64b|plan_|tri_|dispatch_|combine_|requant|quantize' -c 800 --csv --log-file gpurun_out/${TAG}_launches_$wl.csv $CMD > gpurun_out/ncu_l_$wl.log 2>&1; echo "launches $wl rc=$?"
    done
    python tools/summarize_ncu.py $TAG > gpurun_out/${TAG}_summarize.log 2>&1
    mkdir -p gpurun_out/${TAG}_profiles && cp profiles/${TAG}_* gpurun_out/${TAG}_profiles/ 2>/dev/null
    du -sh gpurun_out
}

# 1-GPU lease: the driver's two commands as the driver runs them (default flags), with their wall time
r3g() {
    export PYTHONDONTWRITEBYTECODE=1
    t0=$SECONDS; timeout 900 python bench.py > gpurun_out/r3g_bench_n1.json 2> gpurun_out/r3g_bench_n1.err; echo "default bench rc=$? wall $((SECONDS-t0)) s"
    t0=$SECONDS; timeout 600 python bench.py --impl reference > gpurun_out/r3g_bench_reference.json 2> gpurun_out/r3g_bench_reference.err; echo "reference arm rc=$? wall $((SECONDS-t0)) s"
    tail -c 300 gpurun_out/r3g_bench_n1.json
}

# 1-GPU lease: three-accumulator fused down projection with pipelined TMEM loads in its epilogues — parity, then lines
r3h() {
    timeout 900 python -m pytest tests/test_gpu_tri.py tests/test_gpu_fp8.py -m gpu -q --timeout 600 > gpurun_out/r3h_pytest.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/r3h_pytest.log
    for wl in prefill_w8a8 prefill_fp8; do
    timeout 300 python bench.py --workload $wl --steps 20 --warmup 3 --no-cpu-baseline --no-workloads > gpurun_out/r3h_bench_$wl.json 2> gpurun_out/r3h_bench.err
    echo "$wl exit $? $(python -c "import json; d=json.loads(open('gpurun_out/r3h_bench_$wl.json').read().strip().splitlines()[-1]); print(round(d['value']/1e6,2), d['ms_per_step'], {k:(v['us_per_launch'] if isinstance(v,dict) else v) for k,v in d['stages'].items()})")"
    done
    timeout 200 python tools/tri_sweep.py --pw 0.95 --ns 65536,16384 > gpurun_out/r3h_tri_sweep_pw95.txt 2>> gpurun_out/r3h_bench.err; cat gpurun_out/r3h_tri_sweep_pw95.txt
    tail -3 gpurun_out/r3h_bench.err
}

# 1-GPU lease: what the driver runs at round end, on the final tree: smoke(), pytest -m gpu, bench (its flags), reference arm
r3i() {
    export PYTHONDONTWRITEBYTECODE=1
    timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r3i_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/r3i_smoke.log
    timeout 1500 python -m pytest tests/ -x -q -m gpu > gpurun_out/r3i_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/r3i_pytest.log
    t0=$SECONDS; timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r3i_bench_n1.json 2> gpurun_out/r3i_bench_n1.err; echo "bench rc=$? wall $((SECONDS-t0)) s"
    t0=$SECONDS; timeout 600 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/r3i_bench_reference.json 2> gpurun_out/r3i_bench_reference.err; echo "reference arm rc=$? wall $((SECONDS-t0)) s"
    python - <<PY
import json
d=json.loads(open('gpurun_out/r3i_bench_n1.json').read().strip().splitlines()[-1])
r=json.loads(open('gpurun_out/r3i_bench_reference.json').read().strip().splitlines()[-1])
print('value', round(d['value']/1e6,2), 'e2e', round(d['e2e']['value']/1e6,2), 'roofline', d['roofline']['frac'], 'block', d['block'], 'clocks', d['clocks'])
print('cpu_baseline', d['cpu_baseline'])
print('reference arm', r['value'], r['steps'], r['ms_per_step'], r['cpu_baseline']['sample'])
print({k: round(w['value']/1e6,2) for k,w in d['workloads'].items()})
PY
}

# 1-GPU lease: ncu evidence of the FINAL tree (after the FP8 re-quantiser fix and the fused plan launch): launch lists and
# --set full summaries for the headline, config 3 (W8A8), FP8 and decode.  Summarised on the box (gpurun returns <= 64 MiB).
r3j() {
    export PYTHONDONTWRITEBYTECODE=1
    TAG=r3j
    for wl in ep64k_bf16 prefill_w8a8 prefill_fp8 prefill_bf16 decode_bf16; do
    CMD="python bench.py --workload $wl --steps 2 --warmup 3 --no-cpu-baseline --no-stages --no-workloads"
    $CMD > gpurun_out/${TAG}_plain_$wl.log 2>&1 && \
    timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'i64b|plan_|tri_|dispatch_|combine_|requant|quantize' -c 800 --csv --log-file gpurun_out/${TAG}_launches_$wl.csv $CMD > gpurun_out/ncu_l_$wl.log 2>&1; echo "launches $wl rc=$?"
    done
    CMD="python bench.py --workload ep64k_bf16 --steps 2 --warmup 3 --no-cpu-baseline --no-stages --no-workloads"
    timeout 600 ncu --set full --clock-control none --import-source on -k regex:'i64b_gemm|plan_|dispatch_kernel' -s 16 -c 8 -f -o gpurun_out/${TAG}_full_ep64k_bf16 $CMD > gpurun_out/ncu_f0.log 2>&1; echo "full ep64k rc=$?"
    CMD="python bench.py --workload prefill_w8a8 --steps 2 --warmup 3 --no-cpu-baseline --no-stages --no-workloads"
    timeout 600 ncu --set full --clock-control none --import-source on -k regex:'i64b_gemm|requant|dispatch_kernel|plan_' -s 20 -c 10 -f -o gpurun_out/${TAG}_full_prefill_w8a8 $CMD > gpurun_out/ncu_f1.log 2>&1; echo "full w8a8 rc=$?"
    CMD="python bench.py --workload prefill_fp8 --steps 2 --warmup 3 --no-cpu-baseline --no-stages --no-workloads"
    timeout 600 ncu --set full --clock-control none --import-source on -k regex:'i64b_gemm|requant|dispatch_kernel|plan_' -s 20 -c 10 -f -o gpurun_out/${TAG}_full_prefill_fp8 $CMD > gpurun_out/ncu_f2.log 2>&1; echo "full fp8 rc=$?"
    python tools/summarize_ncu.py $TAG > gpurun_out/${TAG}_summarize.log 2>&1; echo "summarize rc=$?"
    mkdir -p gpurun_out/${TAG}_profiles && cp profiles/${TAG}_* profiles/ncu_traffic.json gpurun_out/${TAG}_profiles/ 2>/dev/null
    rm -f gpurun_out/*.ncu-rep
    du -sh gpurun_out
}

# 1-GPU lease: streaming end-to-end steps (HostStepPipeline.run_stream) — parity, then e2e streamed vs isolated
r3k() {
    timeout 600 python -m pytest tests/test_gpu_host_pipeline.py -m gpu -q --timeout 300 > gpurun_out/r3k_pytest.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/r3k_pytest.log
    for flag in "" "--e2e-isolated" "--e2e-chunks 4" "--e2e-chunks 2"; do
    tag=$(echo "$flag" | tr -d ' -')
    timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-workloads --no-stages $flag > gpurun_out/r3k_bench_$tag.json 2> gpurun_out/r3k_bench.err
    echo "flag='$flag' exit $? $(python -c "import json; d=json.loads(open('gpurun_out/r3k_bench_$tag.json').read().strip().splitlines()[-1]); print(round(d['value']/1e6,2), round(d['e2e']['value']/1e6,2), d['e2e']['ms_per_step'], d['e2e']['copies_per_step'])")"
    done
    for wl in prefill_w8a8 decode_bf16; do
    timeout 300 python bench.py --workload $wl --steps 20 --warmup 5 --no-cpu-baseline --no-workloads --no-stages > gpurun_out/r3k_bench_$wl.json 2>> gpurun_out/r3k_bench.err
    echo "$wl exit $? $(python -c "import json; d=json.loads(open('gpurun_out/r3k_bench_$wl.json').read().strip().splitlines()[-1]); print(round(d['value']/1e6,2), round(d['e2e']['value']/1e6,2), d['e2e']['ms_per_step'])")"
    done
    tail -3 gpurun_out/r3k_bench.err
}

# 1-GPU lease: streamed end-to-end steps, number of micro-batches per step (1 = whole step per slot)
r3l() {
    for wl in ep64k_bf16 prefill_w8a8 decode_bf16; do
    for ch in 1 2 3; do
    timeout 300 python bench.py --workload $wl --steps 20 --warmup 5 --no-cpu-baseline --no-workloads --no-stages --e2e-chunks $ch > gpurun_out/r3l_bench_${wl}_ch$ch.json 2> gpurun_out/r3l_bench.err
    echo "$wl chunks=$ch exit $? $(python -c "import json; d=json.loads(open('gpurun_out/r3l_bench_${wl}_ch$ch.json').read().strip().splitlines()[-1]); print(round(d['value']/1e6,2), round(d['e2e']['value']/1e6,2), d['e2e']['ms_per_step'])")"
    done; done 2>&1 | tee gpurun_out/r3l_e2e_stream_chunks.txt
    tail -2 gpurun_out/r3l_bench.err
}

case "$1" in
    r2e|r2f|r2g|r2h|r2j|r2k|r2l|r2m|r2n|r2o|r2p|r2q|r2r|r2s|r2t|r2u|r2v|r2w|r2x|r2y|r2z|r3a|r3b|r3c|r3d|r3e|r3f|r3g|r3h|r3i|r3j|r3k|r3l) tag=$1; shift; $tag "$@" ;;
    *) echo "usage: $0 {r2e|r2f|r2g|r2h|r2j|r2k|r2l|r2m|r2n|r2o|r2p|r2q|r2r|r2s|r2t|r2u|r2v|r2w|r2x|r2y|r2z|r3a|r3b|r3c|r3d|r3e|r3f|r3g|r3h|r3i|r3j|r3k|r3l} [args]"; exit 2 ;;
esac
