"""The drop-in claim, executed: the REFERENCE's own model class and loader running on top of this repo's modules.

`north_star`: "the engine, scheduler, paged KV cache and OpenAI server run unchanged on top of it".  With this repo's package
directory ahead of the reference's on the path, ``vllm_i64`` overlays itself on the reference package
(vllm_i64/_overlay.py): the reference's ``models/complexity_deep/model.py`` then imports ``TokenRoutedMLP`` /
``DenseSwiGLUMLP`` siblings from this repo, ``core/loader.load_checkpoint`` fills them through the same state-dict keys, and
the logits must match the all-reference model within the bf16 bar.  The reference copy comes from tools/install_ref.py
(oracle/_ref, made by build() in the container; skipped where it is absent)."""
import json
import os
import subprocess
import sys
import tempfile

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "vllm-i64_b200")
REF = os.path.join(ROOT, "oracle", "_ref")
WORKER = os.path.join(ROOT, "tests", "dropin_worker.py")

needs_ref = pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "vllm_i64")),
                               reason="oracle/_ref missing (python tools/install_ref.py in the build container)")


def _run(mode, workdir, device, dtype, path):
    env = dict(os.environ, PYTHONPATH=os.pathsep.join(path), PYTHONDONTWRITEBYTECODE="1")
    env.pop("VLLM_I64_REFERENCE", None)
    p = subprocess.run([sys.executable, WORKER, mode, workdir, device, dtype], env=env, capture_output=True, text=True,
                       timeout=600, cwd=workdir)
    assert p.returncode == 0, p.stderr[-3000:]
    line = [ln for ln in p.stdout.splitlines() if ln.startswith("DROPIN_JSON ")][-1]
    return json.loads(line[len("DROPIN_JSON "):])


@needs_ref
def test_overlay_resolves_reference_model_over_new_layers():
    """CPU: structure only (the new modules have no CPU compute path): same classes' names, same state-dict keys, the
    reference loader fills every parameter."""
    with tempfile.TemporaryDirectory() as wd:
        ref = _run("ref", wd, "none", "fp32", [REF])
        ov = _run("overlay", wd, "none", "fp32", [PKG, REF])
    assert ref["trm_file"].startswith(os.path.realpath(REF))
    assert ov["trm_file"].startswith(os.path.realpath(PKG)), ov["trm_file"]
    assert ov["moe_mlp_file"].startswith(os.path.realpath(PKG))
    assert ov["moe_mlp_class"] == ref["moe_mlp_class"] == "vllm_i64.layers.token_routed_mlp.TokenRoutedMLP"
    assert ov["dense_mlp_file"].startswith(os.path.realpath(PKG)), ov["dense_mlp_file"]
    assert ov["moe_keys"] == ref["moe_keys"] and ov["dense_keys"] == ref["dense_keys"]
    assert ov["moe_load_stats"]["loaded"] > 0


@needs_ref
@pytest.mark.gpu
@pytest.mark.parametrize("dtype,tol_max,tol_mean", [("fp32", 1e-2, 1e-3), ("bf16", 3e-2, 1e-2)])
def test_reference_model_logits_match_on_gpu(dtype, tol_max, tol_mean):
    """GPU: logits of the reference model built over the sm_100a MLPs (loaded by the reference's load_checkpoint) against
    the all-reference model on the same checkpoint and batch, same device and dtype.  fp32: the new block computes in bf16
    tensor-core precision with fp32 accumulation (≤1e-2 / ≤1e-3 on the logits); bf16: both sides round through bf16, the
    bar is the bf16 noise of a 2-layer model."""
    if not torch.cuda.is_available():
        pytest.skip("CUDA device required")
    with tempfile.TemporaryDirectory() as wd:
        _run("ref", wd, "cuda:0", dtype, [REF])
        ov = _run("overlay", wd, "cuda:0", dtype, [PKG, REF])
        assert ov["launches"] > 0, "the overlay model did not launch any kernel of libi64b.so"
        for kind in ("moe", "dense"):
            a = torch.load(os.path.join(wd, f"{kind}_ref_logits.pt"))
            b = torch.load(os.path.join(wd, f"{kind}_overlay_logits.pt"))
            d = (a - b).abs()
            mx, mean = float(d.max() / a.abs().max()), float(d.mean() / a.abs().mean())
            assert mx <= tol_max and mean <= tol_mean, (kind, dtype, mx, mean)


@needs_ref
@pytest.mark.gpu
def test_engine_wiring_on_gpu():
    """SURVEY §8 f2 / f3, executed with the reference's own files: its post-load quantisation passes
    (core/loader.py:627-751, 958-987) switch the new modules to the W8A8 / FP8 kernels (the TokenRoutedMLP pass is dead code
    in the reference: hooked by patch_reference_models()), its CUDAGraphRunner (core/cuda_graph.py:21-153) captures and
    replays the block for every common decode batch size, and the per-layer expert histogram is read from the device."""
    if not torch.cuda.is_available():
        pytest.skip("CUDA device required")
    with tempfile.TemporaryDirectory() as wd:
        _run("ref", wd, "none", "fp32", [REF])          # writes the checkpoints
        env = dict(os.environ, PYTHONPATH=os.pathsep.join([PKG, REF]), PYTHONDONTWRITEBYTECODE="1")
        p = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "wiring_worker.py"), wd], env=env, capture_output=True,
                           text=True, timeout=600, cwd=wd)
        assert p.returncode == 0, p.stderr[-3000:]
        info = json.loads([ln for ln in p.stdout.splitlines() if ln.startswith("WIRING_JSON ")][-1][len("WIRING_JSON "):])
    assert "vllm_i64.core.loader._quantize_experts" in info["hooked"]
    assert info["moe_quantization"] == ["int8"] and info["moe_block_modes"] == [1] and info["moe_int8_launches"] > 0
    assert info["moe_int8_vs_bf16"][1] <= 0.1, info["moe_int8_vs_bf16"]          # int8 noise on a 2-layer model's logits
    assert info["expert_load_layers"] == 2 and info["expert_load_row_sums"] == [48]   # 2 x 24 tokens on every route
    for method in ("int8", "fp8"):
        assert info[f"dense_{method}_modules"] >= 1 and info[f"dense_{method}_quantization"] == [method]
        assert info[f"dense_{method}_vs_bf16"][1] <= (0.1 if method == "int8" else 0.25), (method, info[f"dense_{method}_vs_bf16"])
    assert info["graph_sizes"] == [1, 2, 4, 8, 16, 32, 64]
    assert all(info["graph_replay_equals_eager"]), info["graph_replay_equals_eager"]


@needs_ref
def test_loader_hooks_switch_the_modules_on_cpu():
    """CPU, structure only: with the overlay active, the reference loader's post-load passes (its own file) flip the new
    modules' quantisation mode — `_quantize_experts(model, "int8")` reaches TokenRoutedMLP (dead code in the reference),
    `_quantize_dense_mlp` / `_quantize_dense_mlp_fp8` leave buffers DenseMLP recognises; host-side weights are quantised by
    the reference's own torch ops."""
    code = r'''
import json, torch, torch.nn as nn, vllm_i64
hooked = vllm_i64.patch_reference_models()
from vllm_i64.core import loader
from vllm_i64.layers.dense_mlp import DenseMLP
from vllm_i64.layers.token_routed_mlp import TokenRoutedMLP
class M(nn.Module):
    def __init__(s):
        super().__init__()
        s.a = TokenRoutedMLP(64, 128, 4, 100, shared_expert=True, shared_intermediate_size=32, top_k=2)
        s.d = DenseMLP(64, 128)
m = M(); loader._quantize_experts(m, "int8"); loader._quantize_dense_mlp(m, "int8")
m2 = M(); loader._quantize_dense_mlp_fp8(m2)
print("HOOKS_JSON " + json.dumps({"hooked": hooked, "moe_int8": m.a.quantization, "dense_int8_buf": str(m.d.gate_up_int8.dtype),
      "moe_fp8": m2.a.quantization, "dense_fp8_buf": str(m2.d.gate_fp8.dtype), "load": vllm_i64.collect_expert_load(m)}))
'''
    env = dict(os.environ, PYTHONPATH=os.pathsep.join([PKG, REF]), PYTHONDONTWRITEBYTECODE="1")
    env.pop("VLLM_I64_REFERENCE", None)
    p = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stderr[-3000:]
    info = json.loads([ln for ln in p.stdout.splitlines() if ln.startswith("HOOKS_JSON ")][-1][len("HOOKS_JSON "):])
    assert "vllm_i64.core.loader._quantize_experts" in info["hooked"] and "vllm_i64.core.loader._quantize_dense_mlp_fp8" in info["hooked"]
    assert info["moe_int8"] == "int8" and info["dense_int8_buf"] == "torch.int8"
    assert info["moe_fp8"] == "fp8" and info["dense_fp8_buf"] == "torch.float8_e4m3fn"
    assert info["load"] == {}          # no forward has run: nothing to report


def test_host_pipeline_cuts_are_a_partition():
    """HostStepPipeline's micro-batch boundaries (pure host arithmetic): a partition of [0, n), inner cuts on multiples of
    256, proportional to the weights, never larger than the buffers sized for the full step."""
    sys.path.insert(0, PKG)
    from vllm_i64.kernels.block import HostStepPipeline as H
    for weights in ([1.0] * 4, [1, 2, 3, 4, 3, 2, 1], [1], [5, 1]):
        p = H.__new__(H)
        p.weights, p.chunks = [float(w) for w in weights], len(weights)
        full = p._cuts(65536)
        per = max(hi - lo for lo, hi in full) + 256
        for n in (0, 1, 255, 256, 700, 1537, 16384, 40000, 65536):
            cuts = p._cuts(n)
            assert len(cuts) == len(weights) and cuts[0][0] == 0 and cuts[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(cuts, cuts[1:])) and all(lo <= hi for lo, hi in cuts)
            assert all(hi % 256 == 0 or hi == n for _, hi in cuts[:-1])
            assert max(hi - lo for lo, hi in cuts) <= per
    p = H.__new__(H)
    p.weights, p.chunks = [1.0, 2.0, 3.0, 4.0, 3.0, 2.0, 1.0], 7
    sizes = [hi - lo for lo, hi in p._cuts(65536)]
    assert sizes == sorted(sizes[:4]) + sorted(sizes[4:], reverse=True) and sizes[3] == max(sizes)
