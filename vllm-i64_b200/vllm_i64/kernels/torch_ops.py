"""``torch.ops.vllm_i64.*`` — the op registration of csrc/i64_ops_binding.cu:400-407 over this build's C ABI.

The reference registers six ops with ``TORCH_LIBRARY(vllm_i64, m)`` (route_tokens, silu_hadamard, scatter_by_expert,
gather_by_expert, atomic_scatter, fused_route_scatter) and wraps its whole routed layer in one opaque custom op with a
fake impl so that ``torch.compile`` traces through it (layers/token_routed_i64/fused_i64_moe.py:123-142;
``--compile``, cli.py:500).  Same surface here, registered through ``torch.library`` (the C ABI has no torch types in
it, so the shim lives on the Python side of the boundary):

  * the six ops, reference names and argument order, CUDA implementations = ``kernel_loader.I64Ops`` (one C-ABI call
    each), fake impls giving the output shapes / dtypes;
  * ``add_rmsnorm``            — the fused residual-add + RMSNorm (+ INT8 quant) in front of the block:
    ``(h, y, q, scale)``; ``h`` is empty without a residual, ``q`` / ``scale`` are empty without ``quantize``;
  * ``token_routed_mlp``       — one whole ``TokenRoutedMLP`` / ``DenseMLP`` forward as ONE graph node (the layer is looked
    up by an integer id, so the op's arguments are tensors and ints only); ``forward()`` of those modules routes
    through it while ``torch.compile`` is tracing, so a compiled model has no graph break at the block.

There is no CPU kernel behind any of these ops: calling one with CPU tensors raises (NotImplementedError from the
dispatcher).  The library fragment is created once per process (``FRAGMENT``: coexists with a reference build that
already owns the namespace).
"""
from __future__ import annotations

import weakref
from typing import Optional, Tuple

import torch

_LIB = None
_LAYERS: "weakref.WeakValueDictionary[int, torch.nn.Module]" = weakref.WeakValueDictionary()
_NEXT_ID = [1]

SCHEMAS = {
    # name: schema — the first six are csrc/i64_ops_binding.cu:32, 74, 179, 207, 220, 302
    "route_tokens": "(Tensor token_ids, int num_experts) -> Tensor",
    "silu_hadamard": "(Tensor gate_up, int expert_inter) -> Tensor",
    "scatter_by_expert": "(Tensor hidden, Tensor expert_ids, int num_experts) -> (Tensor, Tensor, Tensor, Tensor)",
    "gather_by_expert": "(Tensor expert_out, Tensor scatter_indices) -> Tensor",
    "atomic_scatter": "(Tensor hidden, Tensor expert_ids, int num_experts) -> (Tensor, Tensor, Tensor, Tensor)",
    "fused_route_scatter": "(Tensor token_ids, Tensor hidden, int num_experts) -> (Tensor, Tensor, Tensor, Tensor, Tensor)",
    "add_rmsnorm": "(Tensor x, Tensor? residual, Tensor weight, float eps, bool quantize) -> (Tensor, Tensor, Tensor, Tensor)",
    "token_routed_mlp": "(Tensor x, Tensor? token_ids, Tensor? x_q, Tensor? x_scale, int layer_id) -> Tensor",
}
REFERENCE_OPS = ("route_tokens", "silu_hadamard", "scatter_by_expert", "gather_by_expert", "atomic_scatter",
                 "fused_route_scatter")


def register_layer(module: torch.nn.Module) -> int:
    """Integer handle under which ``torch.ops.vllm_i64.token_routed_mlp`` finds a module (weakly held)."""
    lid = getattr(module, "_i64_layer_id", None)
    if lid is None:
        lid = _NEXT_ID[0]
        _NEXT_ID[0] += 1
        object.__setattr__(module, "_i64_layer_id", lid)
    _LAYERS[lid] = module
    return lid


class TracedLayer:
    """Mixin for modules that ``vllm_i64::token_routed_mlp`` dispatches to: a copy (``copy.deepcopy`` / unpickling goes
    through ``__setstate__``) gets its own id instead of aliasing the original's."""

    def __setstate__(self, state):
        state = dict(state)
        state.pop("_i64_layer_id", None)
        super().__setstate__(state)
        register_layer(self)


def _ops():
    from vllm_i64.kernels.kernel_loader import get_ops
    return get_ops()


# ------------------------------------------------------------------------------------------------ CUDA implementations
def _route_tokens(token_ids, num_experts):
    return _ops().route_tokens(token_ids, num_experts)


def _silu_hadamard(gate_up, expert_inter):
    return _ops().silu_hadamard(gate_up, expert_inter)


def _scatter_by_expert(hidden, expert_ids, num_experts):
    return _ops().scatter_by_expert(hidden, expert_ids, num_experts)


def _gather_by_expert(expert_out, scatter_indices):
    return _ops().gather_by_expert(expert_out, scatter_indices)


def _atomic_scatter(hidden, expert_ids, num_experts):
    return _ops().atomic_scatter(hidden, expert_ids, num_experts)


def _fused_route_scatter(token_ids, hidden, num_experts):
    return _ops().fused_route_scatter(token_ids, hidden, num_experts)


def _add_rmsnorm(x, residual, weight, eps, quantize):
    from vllm_i64.kernels.cuda import get_i64_cuda_ops
    h, y, qs = get_i64_cuda_ops().add_rmsnorm(x, residual, weight, eps, quantize=quantize, want_normed=True)
    if h is None:                        # no residual: nothing was added (an op may not return its input)
        h = x.new_empty(0)
    if qs is None:
        q = torch.empty(0, dtype=torch.int8, device=x.device)
        s = torch.empty(0, dtype=torch.float32, device=x.device)
    else:
        q, s = qs
    return h, y, q, s


def _token_routed_mlp(x, token_ids, x_q, x_scale, layer_id):
    layer = _LAYERS.get(int(layer_id))
    if layer is None:
        raise RuntimeError(f"vllm_i64::token_routed_mlp: no live layer with id {layer_id}")
    preq = (x_q, x_scale) if x_q is not None else None
    return layer._forward_eager(x, token_ids, preq)


# ------------------------------------------------------------------------------------------------ fake (meta) implementations
def _fake_route_tokens(token_ids, num_experts):
    return torch.empty(token_ids.shape, dtype=torch.int32, device=token_ids.device)


def _fake_silu_hadamard(gate_up, expert_inter):
    return gate_up.new_empty(gate_up.shape[0], expert_inter)


def _fake_scatter(hidden, expert_ids, num_experts):
    i32 = dict(dtype=torch.int32, device=hidden.device)
    return (torch.empty_like(hidden), torch.empty(hidden.shape[0], **i32), torch.empty(num_experts, **i32),
            torch.empty(num_experts, **i32))


def _fake_gather(expert_out, scatter_indices):
    return expert_out.new_empty(scatter_indices.numel(), expert_out.shape[1])


def _fake_fused_route_scatter(token_ids, hidden, num_experts):
    return (_fake_route_tokens(token_ids, num_experts),) + _fake_scatter(hidden, token_ids, num_experts)


def _fake_add_rmsnorm(x, residual, weight, eps, quantize):
    n = x.numel() // x.shape[-1]
    q = torch.empty(x.shape if quantize else (0,), dtype=torch.int8, device=x.device)
    s = torch.empty((n, 1) if quantize else (0,), dtype=torch.float32, device=x.device)
    return (torch.empty_like(x) if residual is not None else x.new_empty(0)), torch.empty_like(x), q, s


def _fake_token_routed_mlp(x, token_ids, x_q, x_scale, layer_id):
    return torch.empty_like(x)


_IMPLS = {
    "route_tokens": (_route_tokens, _fake_route_tokens),
    "silu_hadamard": (_silu_hadamard, _fake_silu_hadamard),
    "scatter_by_expert": (_scatter_by_expert, _fake_scatter),
    "gather_by_expert": (_gather_by_expert, _fake_gather),
    "atomic_scatter": (_atomic_scatter, _fake_scatter),
    "fused_route_scatter": (_fused_route_scatter, _fake_fused_route_scatter),
    "add_rmsnorm": (_add_rmsnorm, _fake_add_rmsnorm),
    "token_routed_mlp": (_token_routed_mlp, _fake_token_routed_mlp),
}


def register() -> None:
    """Define and implement the ops (idempotent).  Ops a reference build already defined keep its schema and gain this
    build's CUDA implementation only if none is registered."""
    global _LIB
    if _LIB is not None:
        return
    lib = torch.library.Library("vllm_i64", "FRAGMENT")
    for name, schema in SCHEMAS.items():
        exists = hasattr(torch.ops.vllm_i64, name)
        if not exists:
            lib.define(name + schema, tags=(torch.Tag.needs_fixed_stride_order,))
        impl, fake = _IMPLS[name]
        try:
            lib.impl(name, impl, "CUDA")
            torch.library.register_fake(f"vllm_i64::{name}", fake, lib=lib)
        except RuntimeError:
            if not exists:
                raise
    _LIB = lib


register()


def token_routed_mlp(x: torch.Tensor, token_ids: Optional[torch.Tensor], x_preq: Optional[Tuple[torch.Tensor, torch.Tensor]],
                     module: torch.nn.Module) -> torch.Tensor:
    """The traced form of a block forward: one ``vllm_i64::token_routed_mlp`` node."""
    q, s = (x_preq[0], x_preq[1]) if x_preq is not None else (None, None)
    return torch.ops.vllm_i64.token_routed_mlp(x, token_ids, q, s, module._i64_layer_id)
