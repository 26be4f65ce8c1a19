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
