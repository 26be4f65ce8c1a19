#!/usr/bin/env python
"""MoE-FFN throughput benchmark (BASELINE.json metric: MoE-FFN tokens/s at 1/2/4/8 B200 + % of roofline).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload decode_bf16|prefill_w8a8|dense_decode|dense_prefill]
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...
    python bench.py --impl reference ...        # the reference's CPU path (oracle port) on the host cores

One "token" = one row of x through one full MoE FFN block of one layer.  One "step" = one pass of the block over
one batch of N tokens for each of the model's L layers (L distinct weight sets and routing tables, so decode is an
HBM number and not an L2 number).  Default workload = BASELINE.json configs[1]: tr-hash-moe-500m, bf16 decode,
256 concurrent sequences, 1 GPU.  Prints ONE JSON line (rank 0).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "vllm-i64_b200"), os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import numpy as np  # noqa: E402
import torch  # noqa: E402

# INFERRED architecture shapes: the checkpoint config is not in the reference repo.
WORKLOADS = {
    "decode_bf16": dict(desc="tr-hash-moe-500m bf16 decode, 256 concurrent sequences", H=1024, E=8, I=512, S=1280,
                        V=32000, K=2, pw=0.5, rs=2.0, ss=1.0, L=24, N=256, mode="bf16"),
    "prefill_w8a8": dict(desc="tr-hash-moe-500m INT8 W8A8 prefill, 16K-token batch", H=1024, E=8, I=512, S=1280,
                         V=32000, K=2, pw=0.5, rs=2.0, ss=1.0, L=4, N=16384, mode="int8"),
    "prefill_bf16": dict(desc="tr-hash-moe-500m bf16 prefill, 16K-token batch", H=1024, E=8, I=512, S=1280,
                         V=32000, K=2, pw=0.5, rs=2.0, ss=1.0, L=4, N=16384, mode="bf16"),
    "decode_w8a8": dict(desc="tr-hash-moe-500m INT8 W8A8 decode, 256 concurrent sequences", H=1024, E=8, I=512,
                        S=1280, V=32000, K=2, pw=0.5, rs=2.0, ss=1.0, L=24, N=256, mode="int8"),
    "tr_moe_306_decode": dict(desc="tr-moe-306 bf16 decode, 256 concurrent sequences", H=1024, E=4, I=768, S=1024,
                              V=32000, K=2, pw=0.5, rs=1.0, ss=1.0, L=18, N=256, mode="bf16"),
    "dense_decode": dict(desc="dense-306 matched dense FFN, 256 tokens", H=1024, E=0, I=0, S=4096, V=32000, K=0,
                         pw=1.0, rs=1.0, ss=1.0, L=18, N=256, mode="bf16"),
    "dense_prefill": dict(desc="dense-306 matched dense FFN, 16K tokens", H=1024, E=0, I=0, S=4096, V=32000, K=0,
                          pw=1.0, rs=1.0, ss=1.0, L=4, N=16384, mode="bf16"),
}
METRIC, UNIT = "moe_ffn_tokens_per_s", "tokens/s"
_JSON_FD = None


def emit(line):
    """The ONE JSON line of the contract, on the real stdout."""
    text = json.dumps(line) + "\n"
    if _JSON_FD is None:
        sys.stdout.write(text)
        sys.stdout.flush()
    else:
        sys.stdout.flush()
        os.write(_JSON_FD, text.encode())


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return dict(hbm=float(p["hbm_gbs"]), tf_burst=float(p["bf16_tflops"]),
                    tf_sust=float(p.get("bf16_tflops_sustained", p["bf16_tflops"])), src="measured")
    except Exception:
        return dict(hbm=6650.0, tf_burst=1590.0, tf_sust=1400.0, src="fallback")


def ncu_traffic(workload, kernel_key):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant kernel, from the committed ncu --set full
    capture of this workload (profiles/ncu_traffic.json, written by tools/summarize_ncu.py); None if not captured."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            t = json.load(f).get(workload, {})
        for name, v in t.items():
            if kernel_key in name:
                return int(v)
    except Exception:
        pass
    return None


def route_weights(wl):
    k = wl["K"]
    if k <= 1:
        return [1.0]
    return [wl["pw"]] + [(1.0 - wl["pw"]) / (k - 1)] * (k - 1)


# ------------------------------------------------------------------------------------------------- algorithmic work
def stage_bytes(wl):
    """ALGORITHMIC HBM bytes per launch of each stage for one layer, and flops of the GEMMs."""
    H, E, I, S, K, N = (wl[k] for k in "HEISKN")
    a = 1 if wl["mode"] == "int8" else 2               # bytes per operand element fed to the tensor cores
    R = N * K
    has_s = 1 if S else 0
    rows = R + N * has_s
    b = {}
    b["plan"] = N * 8 + R * 4
    drows = rows if a == 1 else R          # bf16: the shared expert reads x in place (no token-order copy)
    b["dispatch"] = (N * H * 2 + drows * (H * a + (4 if a == 1 else 0)) + R * 4 * 3) if drows else 0
    w13 = (E * 2 * I + 2 * S) * H * a
    inter_w = (R * I + N * S) * (4 if a == 1 else 2)   # W8A8: Q14 int32 out; bf16: bf16 out
    b["gemm1"] = w13 + rows * H * a + inter_w
    b["requant"] = (R * I + N * S) * 5 if a == 1 else 0
    w2 = (E * H * I + H * S) * a
    if has_s and E == 0:
        # fused combine (dense FFN): the shared expert's GEMM2 tiles pull the routed fp32 rows and write the block output
        b["gemm2"] = w2 + (R * I + N * S) * a + 2 * R * H * 4 + R * 4 + N * H * 2
        b["combine"] = 0
    else:
        b["gemm2"] = w2 + (R * I + N * S) * a + rows * H * 4
        b["combine"] = rows * H * 4 + N * H * 2 + R * 4
    f = {"gemm1": 2.0 * H * 2 * (R * I + N * S), "gemm2": 2.0 * H * (R * I + N * S)}
    return b, f


# ------------------------------------------------------------------------------------------------- clocks sampling
class ClockSampler:
    def __init__(self, index):
        self.samples, self.reasons, self._stop, self.max_mhz = [], set(), threading.Event(), None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None
        self.t = None

    def _loop(self):
        nv = self.nv
        names = {nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
                 nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap"}
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.004)

    def __enter__(self):
        if self.nv:
            self.t = threading.Thread(target=self._loop, daemon=True)
            self.t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self.t:
            self.t.join()

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": 0}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


# ------------------------------------------------------------------------------------------------- model build
def build_layers(wl, device):
    """L blocks with distinct random-init weights (reference init range) and layer-specific hash tables."""
    from vllm_i64 import _C
    from vllm_i64.kernels.block import I64Block
    from vllm_i64.core.quantization import quantize_int8
    from tools.synth_data import synthetic_hash_table
    H, E, I, S, V, K = (wl[k] for k in "HEISVK")
    mode = _C.MODE_W8A8 if wl["mode"] == "int8" else _C.MODE_BF16
    layers = []
    for layer in range(wl["L"]):
        g = torch.Generator(device=device).manual_seed(layer)

        def u(shape, fan_in):
            b = 1.0 / fan_in ** 0.5
            return (torch.rand(shape, device=device, generator=g) * 2 - 1) * b

        blk = I64Block(H, E, I, S, V, K, route_weights(wl), wl["rs"], wl["ss"], mode, device)
        gate = u((E, H, I), H) if E else None
        up = u((E, H, I), H) if E else None
        down = u((E, I, H), I) if E else None
        sg, su, sd = (u((S, H), H), u((S, H), H), u((H, S), S)) if S else (None, None, None)
        if mode == _C.MODE_BF16:
            blk.load_bf16(gate, up, down, sg, su, sd)
        else:
            ex = [None] * 4
            if E:
                q13, s13 = quantize_int8(torch.cat([gate.transpose(1, 2), up.transpose(1, 2)], 1).contiguous())
                q2, s2 = quantize_int8(down.transpose(1, 2).contiguous())
                ex = [q13, s13, q2, s2]
            sh = [None] * 4
            if S:
                q, s_ = quantize_int8(torch.cat([sg, su], 0).contiguous())
                q2s, s2s = quantize_int8(sd.contiguous())
                sh = [q, s_, q2s, s2s]
            blk.load_w8a8(*ex, *sh)
        if E:
            blk.set_table(torch.from_numpy(synthetic_hash_table(V, E, K, layer)))
        layers.append(blk)
    torch.cuda.synchronize(device)
    for b in layers:
        b._keep = None
    # decoder order: layer l is followed by l+1 (the step loop wraps around), so each block may prefetch the next one's
    # weights into L2 (i64b_moe_set_next; a no-op outside the fused decode kernel).  Opt-in: measured slower
    if len(layers) > 1 and os.environ.get("I64B_PREFETCH", "0") != "0":
        for i, b in enumerate(layers):
            b.set_next(layers[(i + 1) % len(layers)])
    return layers


def time_graph(run, steps, warmup, stream, barrier=None, sampler=None):
    """Capture `run` once, replay `warmup` then `steps` times; CUDA events on the launching stream."""
    with torch.cuda.stream(stream):
        run()                                   # eager warm (workspace growth, lazy init) before capture
    stream.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=stream):
        run()
    with torch.cuda.stream(stream):
        for _ in range(max(warmup, 3)):
            g.replay()
        stream.synchronize()
        if barrier:
            barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ctx = sampler if sampler is not None else _Null()
        with ctx:
            e0.record(stream)
            for _ in range(steps):
                g.replay()
            e1.record(stream)
            e1.synchronize()
        torch.cuda.synchronize()
        if barrier:
            barrier()
    return e0.elapsed_time(e1)


class _Null:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False


# ------------------------------------------------------------------------------------------------- CPU arm
def cpu_reference_run(wl, steps, warmup, mode="int8", max_seconds=25.0, layers_cap=None):
    """Reference CPU path (oracle port, oracle/cpu_reference.py) on the host cores: tokens/s over bounded steps."""
    from oracle.cpu_reference import CpuTokenRoutedMLP, quantize_cpu_dynamic, set_cpu_threads
    from tools.synth_data import synthetic_hash_table
    threads = set_cpu_threads()
    H, E, I, S, V, K, N = (wl[k] for k in "HEISVKN")
    if E == 0:      # dense FFN: one "expert" that every token uses
        E_, I_, S_, K_ = 1, S, 0, 1
    else:
        E_, I_, S_, K_ = E, I, S, K
    L = min(wl["L"], layers_cap or wl["L"])
    mods = []
    g = torch.Generator().manual_seed(0)
    ids = torch.randint(0, V, (N,), generator=g)
    x = torch.randn(N, H, generator=g)
    for layer in range(L):
        torch.manual_seed(layer)
        table = torch.from_numpy(synthetic_hash_table(V, E_, K_, layer)) if E_ > 1 else torch.zeros(K_, V, dtype=torch.long)
        m = CpuTokenRoutedMLP(H, E_, I_, S_, K_, wl["pw"], wl["rs"], wl["ss"], table)
        mods.append(quantize_cpu_dynamic(m) if mode == "int8" else m.eval())

    def step():
        with torch.no_grad():
            for m in mods:
                m(x, ids)

    for _ in range(max(1, min(warmup, 2))):
        step()
    t0 = time.perf_counter()
    done = 0
    for _ in range(max(1, steps)):
        step()
        done += 1
        if time.perf_counter() - t0 > max_seconds:
            break
    dt = time.perf_counter() - t0
    return dict(value=done * N * L / dt, ms_per_step=dt / done * 1e3, steps=done, threads=threads, layers=L,
                sample=f"{done} step(s) x {L} layer(s) x {N} tokens, {'dynamic INT8 (x86/FBGEMM)' if mode == 'int8' else 'fp32'}"
                       f", {threads} threads")


# ------------------------------------------------------------------------------------------------- expert parallel
def run_ep(wl, device, rank, world, args):
    """BASELINE config 5: experts sharded over the ranks, 64K tokens per step in total (strong scaling), NCCL
    all-to-all over NVLink for the token exchange.  Timed on the device, max over ranks."""
    import torch.distributed as dist
    from vllm_i64 import _C
    from vllm_i64.parallel.expert_parallel import EPTokenRoutedBlock, init_expert_parallel
    from tools.synth_data import synthetic_hash_table
    ep = init_expert_parallel()
    H, E, I, S, V, K = (wl[k] for k in "HEISVK")
    L = 4
    n = args.ep_tokens // world
    blocks = []
    tdt = torch.bfloat16 if args.ep_transport == "bf16" else torch.float32
    for layer in range(L):
        g = torch.Generator(device=device).manual_seed(layer)

        def u(shape, fan_in):
            return (torch.rand(shape, device=device, generator=g) * 2 - 1) / fan_in ** 0.5

        b = EPTokenRoutedBlock(H, E, I, S, V, K, route_weights(wl), wl["rs"], wl["ss"], _C.MODE_BF16, device, ep,
                               transport_dtype=tdt)
        b.load_bf16(u((E, H, I), H), u((E, H, I), H), u((E, I, H), I), u((S, H), H), u((S, H), H), u((H, S), S))
        b.set_table(torch.from_numpy(synthetic_hash_table(V, E, K, layer)))
        blocks.append(b)
    x = torch.randn(n, H, device=device, generator=torch.Generator(device=device).manual_seed(100 + rank)).to(torch.bfloat16)
    ids = torch.randint(0, V, (n,), device=device, generator=torch.Generator(device=device).manual_seed(200 + rank))

    outs = [torch.empty(n, H, dtype=torch.bfloat16, device=device) for _ in range(2)]

    def timed(exchange, steps, warm):
        def step():
            for i, b in enumerate(blocks):
                b.forward(x, ids, torch.bfloat16, exchange=exchange, out=outs[i & 1])
        for _ in range(warm):
            step()
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            step()
        e1.record()
        e1.synchronize()
        dist.barrier()
        t = torch.tensor([e0.elapsed_time(e1)], device=device, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0]) / steps

    steps = max(3, min(args.steps, 20))
    ms = timed("peer", steps, 3)                 # product path: NVLink peer-memory exchange fused into the kernels
    ms_nccl = timed("nccl", max(3, steps // 4), 2)   # reference-style all_to_all_single exchange (baseline)
    sent = n * K * (world - 1) / world * H * 2
    return {"workload": "tr-hash-moe-500m bf16 expert-parallel", "tokens_per_step_total": n * world,
            "layers_per_step": L, "ms_per_step": ms, "value": n * world * L / (ms * 1e-3), "unit": UNIT,
            "scaling": "strong", "a2a_bytes_out_per_rank_per_layer": int(sent), "ep_size": world,
            "transport_dtype": args.ep_transport,
            "exchange": "NVLink peer memory (dispatch kernel writes into the owners' buffers; GEMM2 epilogue writes "
                        "outputs home); NCCL only all-gathers the token ids",
            "nccl_all_to_all_baseline": {"ms_per_step": ms_nccl, "value": n * world * L / (ms_nccl * 1e-3)}}


# ------------------------------------------------------------------------------------------------- main
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=300)
    ap.add_argument("--warmup", type=int, default=30)
    ap.add_argument("--workload", default="decode_bf16", choices=sorted(WORKLOADS))
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--cpu-mode", default="int8", choices=["int8", "fp32"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-stages", action="store_true")
    ap.add_argument("--no-ep", action="store_true")
    ap.add_argument("--ep-tokens", type=int, default=65536, help="tokens per step over ALL ranks (BASELINE config 5)")
    ap.add_argument("--ep-transport", default="f32", choices=["f32", "bf16"],
                    help="dtype the expert outputs travel back in (f32: bit-identical to one GPU; bf16: the reference's)")
    args = ap.parse_args()
    wl = dict(WORKLOADS[args.workload])
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    cfg_json = {"workload": wl["desc"], "hidden": wl["H"], "num_experts": wl["E"], "expert_inter": wl["I"],
                "shared_inter": wl["S"], "top_k": wl["K"], "layers_per_step": wl["L"], "tokens_per_gpu": wl["N"],
                "shapes": "INFERRED (SURVEY.md §8)", "parallelism": f"tokens sharded over {world} GPU(s), experts replicated",
                "l2_policy": "L distinct layer weight sets per step (>> 126 MB L2)"}

    if args.impl == "reference":
        if rank != 0:
            return
        r = cpu_reference_run(wl, args.steps, args.warmup, args.cpu_mode)
        line = {"metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": r["steps"],
                "warmup": min(args.warmup, 2), "ms_per_step": r["ms_per_step"], "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "int8" if args.cpu_mode == "int8" else "f32",
                "data": "synthetic", "impl": "reference", "config": cfg_json,
                "cpu_baseline": {"value": r["value"], "unit": UNIT, "cores": r["threads"], "kind": "port",
                                 "sample": r["sample"]},
                "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        emit(line)
        return

    assert torch.cuda.is_available(), "bench.py needs a GPU (use --impl reference for the CPU arm)"
    device = torch.device("cuda", local_rank)
    torch.cuda.set_device(device)
    barrier = None
    if world > 1 or os.environ.get("I64B_BENCH_FORCE_DIST") == "1":      # (the env is a bring-up switch: NCCL at world 1)
        import torch.distributed as dist
        # NCCL prints its version banner to stdout when NCCL_DEBUG is set by the environment: keep stdout to the ONE
        # JSON line of the contract
        # (the banner is written to file descriptor 1 directly, so fd 1 is pointed at stderr for the whole run and the
        # JSON line goes to the saved descriptor)
        global _JSON_FD
        sys.stdout.flush()
        _JSON_FD = os.dup(1)
        os.dup2(2, 1)
        dist.init_process_group("nccl", device_id=device)
        barrier = lambda: dist.barrier()            # noqa: E731
    from vllm_i64 import _C

    H, N, L = wl["H"], wl["N"], wl["L"]
    layers = build_layers(wl, device)
    g = torch.Generator().manual_seed(1234 + rank)
    ids_host = torch.randint(0, wl["V"], (N,), generator=g).pin_memory()
    x_host = torch.randn(N, H, generator=torch.Generator().manual_seed(rank)).to(torch.bfloat16).pin_memory()
    out_host = torch.empty(N, H, dtype=torch.bfloat16).pin_memory()
    x_dev, ids_dev = x_host.to(device), ids_host.to(device)
    outs = [torch.empty(N, H, dtype=torch.bfloat16, device=device) for _ in range(2)]
    ids_arg = ids_dev if wl["E"] else None
    stream = torch.cuda.Stream(device)

    def step_device(stages=_C.STAGE_ALL):
        for i, blk in enumerate(layers):
            blk.forward(x_dev, ids_arg, torch.bfloat16, out=outs[i & 1], stages=stages)

    def step_e2e():
        x_dev.copy_(x_host, non_blocking=True)
        ids_dev.copy_(ids_host, non_blocking=True)
        step_device()
        out_host.copy_(outs[(L - 1) & 1], non_blocking=True)

    # launches per step (counted by the library itself, eager pass)
    with torch.cuda.stream(stream):
        step_device()
        stream.synchronize()
        _C.launch_count_reset()
        step_device()
        stream.synchronize()
    launches_per_step = _C.launch_count()

    sampler = ClockSampler(local_rank)
    ms = time_graph(step_device, args.steps, args.warmup, stream, barrier, sampler)
    ms_e2e = time_graph(step_e2e, args.steps, args.warmup, stream, barrier)
    tokens_step = N * L

    if world > 1:
        import torch.distributed as dist
        t = torch.tensor([ms, ms_e2e], device=device, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, ms_e2e = float(t[0]), float(t[1])
    value = world * tokens_step * args.steps / (ms * 1e-3)
    e2e_value = world * tokens_step * args.steps / (ms_e2e * 1e-3)

    # per-stage timing (isolated replays of each kernel over the L layers) and the roofline of the dominant one
    pk = peaks()
    stages_json, roof = {}, None
    if rank == 0 and not args.no_stages:
        sb, sf = stage_bytes(wl)
        names = [("plan", _C.STAGE_PLAN), ("dispatch", _C.STAGE_DISPATCH), ("gemm1", _C.STAGE_GEMM1),
                 ("requant", _C.STAGE_REQUANT), ("gemm2", _C.STAGE_GEMM2), ("combine", _C.STAGE_COMBINE)]
        reps = max(20, min(args.steps, 100))
        for name, bit in names:
            if sb[name] == 0:
                continue
            t_ms = time_graph(lambda b=bit: step_device(b), reps, 5, stream)
            us = t_ms * 1e3 / (reps * L)
            gbs = sb[name] / (us * 1e-6) / 1e9
            ent = {"us_per_launch": round(us, 3), "algorithmic_bytes": sb[name], "GBps": round(gbs, 1),
                   "frac_hbm": round(gbs / pk["hbm"], 4)}
            if name in sf:
                tf = sf[name] / (us * 1e-6) / 1e12
                ent["TFLOPs" if wl["mode"] == "bf16" else "TOPs"] = round(tf, 1)
                ent["frac_tensor_bf16_peak"] = round(tf / pk["tf_burst"], 4)
            stages_json[name] = ent
        dom = max(stages_json, key=lambda k: stages_json[k]["us_per_launch"])
        d = stages_json[dom]
        hbm_bound = d["frac_hbm"] >= d.get("frac_tensor_bf16_peak", 0.0) / (2.0 if wl["mode"] == "int8" else 1.0)
        kkey = {"gemm1": "i64b_gemm_kernel<%d, 1" % (1 if wl["mode"] == "int8" else 0),
                "gemm2": "i64b_gemm_kernel<%d, 0" % (1 if wl["mode"] == "int8" else 0)}.get(dom, dom)
        if hbm_bound:
            roof = {"kernel": dom, "bound": "hbm", "achieved": d["GBps"], "peak": pk["hbm"], "unit": "GB/s",
                    "frac": d["frac_hbm"], "traffic": None, "peak_source": pk["src"] + " (burst copy)"}
        else:
            key = "TFLOPs" if wl["mode"] == "bf16" else "TOPs"
            peak = pk["tf_burst"] * (2.0 if wl["mode"] == "int8" else 1.0)
            roof = {"kernel": dom, "bound": "tensor", "achieved": d[key], "peak": peak,
                    "unit": "TFLOP/s" if wl["mode"] == "bf16" else "TOP/s", "frac": round(d[key] / peak, 4),
                    "traffic": None, "peak_source": pk["src"] + (" bf16 burst x2 (nominal INT8:bf16 ratio)" if wl["mode"] == "int8" else " bf16 burst")}
        if roof is not None:
            roof["traffic"] = ncu_traffic(args.workload, kkey)
        if layers[0].uses_fused(N):
            # the step runs ONE fused kernel per layer; the stage list above is the unfused chain, for comparison
            stages_json = {"unfused_chain_for_comparison": stages_json}
        sum_us = sum(v["us_per_launch"] for v in stages_json.values() if isinstance(v, dict) and "us_per_launch" in v)
        stages_json["_sum_us_per_layer"] = round(sum_us, 2)
        stages_json["_step_us_per_layer"] = round(ms * 1e3 / (args.steps * L), 2)
        # whole-block HBM roofline: algorithmic minimum bytes of the fused block per layer
        a = 1 if wl["mode"] == "int8" else 2
        min_bytes = ((wl["E"] * 3 * wl["I"] + 3 * wl["S"]) * H * a) + N * H * 2 * 2
        stages_json["_block_min_bytes"] = min_bytes
        stages_json["_block_frac_hbm"] = round(min_bytes / (ms * 1e-3 / (args.steps * L)) / 1e9 / pk["hbm"], 4)

    if rank == 0 and layers[0].uses_fused(N):
        # dominant (only) kernel of the step: the fused layer kernel, timed live over the whole timed region
        us = ms * 1e3 / (args.steps * L)
        a_ = 1 if wl["mode"] == "int8" else 2
        fb = ((wl["E"] * 3 * wl["I"] + 3 * wl["S"]) * H * a_) + N * H * 2 * 2     # weights once + x in + out
        gbs = fb / (us * 1e-6) / 1e9
        roof = {"kernel": "i64b_fused_layer_kernel", "bound": "hbm", "achieved": round(gbs, 1), "peak": pk["hbm"],
                "unit": "GB/s", "frac": round(gbs / pk["hbm"], 4),
                "traffic": ncu_traffic(args.workload, "i64b_fused_layer_kernel"),
                "peak_source": pk["src"] + " (burst copy)", "algorithmic_bytes": fb, "us_per_launch": round(us, 3)}
    ep_json = None
    if world > 1 and wl["E"] and wl["E"] % world == 0 and not args.no_ep:
        try:
            ep_json = run_ep(wl, device, rank, world, args)
        except Exception as exc:                      # the data-parallel line above stands on its own
            ep_json = {"error": f"{type(exc).__name__}: {exc}"[:300]}

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        r = cpu_reference_run(wl, steps=3, warmup=1, mode=args.cpu_mode, max_seconds=20.0,
                              layers_cap=4 if N > 4096 else None)
        cpu = {"value": r["value"], "unit": UNIT, "cores": r["threads"], "kind": "port", "sample": r["sample"]}

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "bf16" if wl["mode"] == "bf16" else "int8",
                "data": "synthetic", "config": cfg_json, "clocks": sampler.summary(),
                "e2e": {"value": e2e_value, "unit": UNIT, "ms_per_step": ms_e2e / args.steps,
                        "h2d_bytes_per_step": N * H * 2 + N * 8, "d2h_bytes_per_step": N * H * 2},
                "gpu_launches": int(launches_per_step) * args.steps, "gpu_launches_per_step": int(launches_per_step),
                "roofline": roof, "stages": stages_json, "cpu_baseline": cpu, "expert_parallel": ep_json}
        emit(line)
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
