"""CPU-side checks of the drop-in boundary: the shared library loads without a GPU, exports every symbol that
include/i64b.h declares, embeds the reference's SiLU table, and the host layer mirrors the reference's interface
(names, arguments, error behaviour).  No compute calls here."""
import ctypes
import inspect
import os
import re

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "i64b.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(i64b_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from vllm_i64 import _C
    lib = _C.lib()
    declared = _declared_symbols()
    assert len(declared) >= 30
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/i64b.h but not exported"
    assert sorted(_C.EXPORTED) == declared, "ctypes signature table out of sync with the header"
    assert lib.i64b_version() >= 100
    assert lib.i64b_last_error() is not None


def test_embedded_silu_lut_matches_reference(golden):
    from vllm_i64 import _C
    p = _C.lib().i64b_silu_lut_q7()
    got = np.ctypeslib.as_array(p, shape=(2049,)).astype(np.int32)
    np.testing.assert_array_equal(got, golden("silu_lut.npy"))


def test_argument_errors_do_not_need_a_gpu():
    from vllm_i64 import _C
    lib = _C.lib()
    cfg = _C.MoEConfig()
    cfg.hidden_size, cfg.num_experts, cfg.expert_inter, cfg.vocab_size, cfg.top_k = 128, 4, 32, 100, 5   # top_k > E
    h = ctypes.c_void_p()
    assert lib.i64b_moe_create(ctypes.byref(cfg), ctypes.byref(h)) == -1
    assert b"top_k" in lib.i64b_last_error()
    cfg.top_k = 2
    assert lib.i64b_moe_create(ctypes.byref(cfg), ctypes.byref(h)) == 0
    assert lib.i64b_moe_packed_bytes(h) > 4 * 3 * 128 * 32 * 2
    assert lib.i64b_moe_workspace_bytes(h, 256) > 256 * 3 * 128 * 2
    # forward before weights are loaded → I64B_ERR_STATE, not a crash
    assert lib.i64b_moe_forward(h, None, None, None, 0, 4, None, 0, None) == -4
    lib.i64b_moe_destroy(h)
    assert lib.i64b_route_modulo(None, None, 10, 6, None) == -1          # E must be a power of two
    assert lib.i64b_plan_workspace_bytes(70000, 2, 8) > 70000 * 2 * 4


def test_host_layer_mirrors_reference_interface():
    from vllm_i64.layers.token_routed_mlp import TokenRoutedMLP
    from vllm_i64.layers.dense_mlp import DenseMLP
    from vllm_i64.kernels import fused_experts, i64_ops, kernel_loader
    from vllm_i64.kernels.cuda import get_i64_cuda_ops
    from vllm_i64.core import quantization
    from vllm_i64.parallel import tensor_parallel

    sig = inspect.signature(TokenRoutedMLP.__init__)
    assert list(sig.parameters)[1:] == [
        "hidden_size", "intermediate_size", "num_experts", "vocab_size", "shared_expert", "shared_intermediate_size",
        "top_k", "top_k_primary_weight", "use_shared_routed_gates", "shared_gate_init", "routed_gate_init",
        "shared_output_scale", "routed_output_scale"]                    # layers/token_routed_mlp.py:24-39
    m = TokenRoutedMLP(128, 128, 4, 512, shared_expert=True, shared_intermediate_size=64, top_k=2,
                       top_k_primary_weight=0.5, use_shared_routed_gates=True)
    assert set(m.state_dict()) == {"gate_proj_w", "up_proj_w", "down_proj_w", "shared_gate.weight", "shared_up.weight",
                                   "shared_down.weight", "shared_output_gate", "routed_output_gate", "token_to_expert",
                                   "topk_token_to_expert"}
    assert m.gate_proj_w.shape == (4, 128, 32) and m.down_proj_w.shape == (4, 32, 128)
    assert torch.equal(m.topk_token_to_expert[1], (m.topk_token_to_expert[0] + 1) % 4)   # tests/test_model.py:198-205
    assert TokenRoutedMLP(128, 128, 4, 512, top_k=2).primary_weight == 0.95              # default, :53-58
    with pytest.raises(ValueError):
        TokenRoutedMLP(128, 128, 4, 512, top_k=5)
    for name in ("route", "expert_forward", "forward", "load_full_weights", "fuse_shared_gate_up",
                 "materialize_expert_linears"):
        assert callable(getattr(m, name))
    full = [torch.randn(4, 128, 32), torch.randn(4, 128, 32), torch.randn(4, 32, 128)]
    m.load_full_weights(*full)
    assert torch.equal(m.gate_proj_w.data, full[0]) and torch.equal(m.down_proj_w.data, full[2])
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        m(torch.randn(3, 128), token_ids=torch.arange(3))
    assert list(inspect.signature(DenseMLP.forward).parameters)[:3] == ["self", "x", "x_preq"]
    ops = kernel_loader.get_ops()
    for name in ("route_tokens", "silu_hadamard", "scatter_by_expert", "gather_by_expert", "atomic_scatter",
                 "fused_route_scatter"):                                   # csrc/i64_ops_binding.cu:400-407
        assert callable(getattr(ops, name))
    cops = get_i64_cuda_ops()
    for name in ("quantize_int8", "dequantize_int8", "quantize_perchannel_int8", "gemm_dequant_int8", "gemm_silu_int8"):
        assert callable(getattr(cops, name))                              # csrc/I64_binding.cu:49-87
    for name in ("i64_route_tokens", "i64_scatter", "i64_gather", "i64_expert_forward", "i64_full_pipeline"):
        assert callable(getattr(i64_ops, name))
    assert list(inspect.signature(fused_experts.fused_token_routed_forward).parameters) == [
        "x", "gate_up_proj", "down_proj", "expert_ids", "num_experts", "intermediate_per_tp", "bmm_threshold"]
    for name in ("quantize_int8", "quantize_activations_int8", "int8_linear_native", "int8_fused_gate_up_native",
                 "dequantize_int8", "quantize_experts"):
        assert callable(getattr(quantization, name))
    for name in ("get_tp", "all_reduce", "shard_expert_weights", "init_distributed"):
        assert callable(getattr(tensor_parallel, name))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ops.route_tokens(torch.arange(8), 4)


def test_dense_int4_buffers_raise_instead_of_running_unquantised():
    """``--quantization int4`` leaves gate_int4 ... buffers on the dense MLP (core/loader.py:685-697).  This build has no INT4
    kernel: it must say so, not run the float weights as if nothing had been asked."""
    from vllm_i64.layers.dense_mlp import DenseMLP
    m = DenseMLP(128, 64)
    m.register_buffer("gate_int4", torch.zeros(64, 64, dtype=torch.uint8))
    with pytest.raises(NotImplementedError, match="INT4"):
        m._ensure_block(torch.device("cpu"))


def test_overlay_int4_functions_refuse_cuda_tensors():
    """Reference functions taken over by the overlay with ``int4`` in their name run on host tensors (load-time packing) and
    raise on CUDA tensors: no PyTorch fallback stands in for a kernel this build does not have."""
    from vllm_i64 import _overlay

    class T:
        def __init__(self, cuda):
            self.is_cuda = cuda

    g = _overlay._guard_int4("kernels.fused_experts.fused_token_routed_forward_int4", lambda x, scale=None: "ran")
    assert g(T(False), scale=T(False)) == "ran"
    with pytest.raises(NotImplementedError, match="int4"):
        g(T(True))
    with pytest.raises(NotImplementedError):
        g(T(False), scale=T(True))
    from vllm_i64.kernels import fused_experts
    if _overlay.load_reference_module("vllm_i64.kernels.fused_experts") is not None:     # overlay active (build container)
        assert fused_experts.fused_token_routed_forward_int4.__wrapped__ is not None


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "vllm-i64_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "oracle" not in text.replace("i64b_", ""), f"{f} references the oracle"


def test_fused_decode_shape_limits():
    """Host arithmetic of the single-launch decode kernel's shape check (no device needed): the named decode shapes
    qualify; shapes whose static work list would overflow the per-CTA unit table (8 units per SM), or that exceed the
    width / batch limits, must take the kernel chain."""
    from vllm_i64 import _C
    ok = _C.lib().i64b_fused_decode_supported
    assert ok(256, 1024, 512, 1280, 8, 2) == 1          # tr-hash-moe-500m decode (config 2)
    assert ok(256, 1024, 768, 1024, 4, 2) == 1          # tr-moe-306
    assert ok(256, 1024, 0, 4096, 0, 0) == 1            # dense-306 (8 slabs, 64 GEMM1 tiles per m-tile)
    assert ok(1024, 256, 0, 1536, 0, 0) == 1            # largest batch
    assert ok(1025, 1024, 512, 1280, 8, 2) == 0         # batch too large
    assert ok(100, 256, 0, 4096 + 64, 0, 0) == 0        # more than 8 split-K slabs
    assert ok(256, 4096, 2048, 0, 64, 2) == 0           # 2048 + 2048 units: more than 8 per SM
    assert ok(256, 1000, 512, 1280, 8, 2) == 0          # hidden not a multiple of 64
