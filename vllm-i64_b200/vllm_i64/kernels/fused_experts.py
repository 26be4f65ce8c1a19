"""``fused_token_routed_forward{,_int8}`` — kernels/fused_experts.py:29-291 on the grouped tcgen05 GEMMs.

The reference switches between a per-token ``bmm`` (N ≤ ``bmm_threshold``, materialising ``[N,H,2I]`` weights)
and a chunked loop with ``.item()`` syncs; both compute the same function, so this build has one path (sort-free
deterministic dispatch → grouped GEMM1+SwiGLU → grouped GEMM2 → un-permute) for every N, CUDA-graph safe.
Packed weights are cached per (tensor storage, version).
"""
from __future__ import annotations

from typing import Dict, Tuple

import torch

from vllm_i64 import _C
from vllm_i64.kernels.block import I64Block

_CACHE: Dict[Tuple, Tuple[I64Block, tuple]] = {}
_CACHE_MAX = 64


def _key(*tensors: torch.Tensor) -> Tuple:
    """What the packed block was made from: the bytes' location (storage address + offset + shape + strides), their type
    and the in-place version.  Views, slices, ``.detach()`` and ``.data`` of a live parameter describe the same bytes and
    give the same key, although each access creates a new Python tensor object."""
    return tuple((t.untyped_storage().data_ptr(), t.storage_offset(), tuple(t.shape), tuple(t.stride()), t.dtype,
                  t._version) for t in tensors)


def _lookup(key: Tuple, tensors) -> "I64Block | None":
    """A hit needs the SAME storages to be alive: an address alone may have been recycled by the allocator after the
    original weights were freed, so the entry holds a weak reference to each storage (the storage object lives as long as
    any tensor uses it, whichever Python object that is)."""
    ent = _CACHE.get(key)
    if ent is None:
        return None
    blk, refs = ent
    if all(r() is t.untyped_storage() for r, t in zip(refs, tensors)):
        return blk
    del _CACHE[key]
    return None


def _store(key: Tuple, tensors, blk: I64Block) -> None:
    import weakref
    if len(_CACHE) >= _CACHE_MAX:
        _CACHE.clear()
    _CACHE[key] = (blk, tuple(weakref.ref(t.untyped_storage()) for t in tensors))


def invalidate_cache() -> None:
    """Drop every packed block.  Call it after rewriting weights in a way the key cannot see: writes through ``.data``,
    and any in-place update of a parameter whose ``.data`` (which has no version counter of its own) is what the caller
    passes to the functions below.  Slices and ``.detach()`` share the parameter's counter and need nothing."""
    _CACHE.clear()


def _cached_block(gate_up_proj: torch.Tensor, down_proj: torch.Tensor, num_experts: int) -> I64Block:
    key = _key(gate_up_proj, down_proj)
    blk = _lookup(key, (gate_up_proj, down_proj))
    if blk is None:
        e, h, two_i = gate_up_proj.shape
        inter = two_i // 2
        if e != num_experts or down_proj.shape != (e, inter, h):
            raise RuntimeError("gate_up_proj must be [E,H,2I] and down_proj [E,I,H]")
        blk = I64Block(h, e, inter, 0, 1, 1, [1.0], 1.0, 1.0, _C.MODE_BF16, gate_up_proj.device)
        blk.load_bf16(gate_up_proj[:, :, :inter], gate_up_proj[:, :, inter:], down_proj)
        _store(key, (gate_up_proj, down_proj), blk)
    return blk


def fused_token_routed_forward(x: torch.Tensor, gate_up_proj: torch.Tensor, down_proj: torch.Tensor,
                               expert_ids: torch.Tensor, num_experts: int, intermediate_per_tp: int,
                               bmm_threshold: int = 64) -> torch.Tensor:
    """``(N,H)`` → ``(N,H)``; ``gate_up_proj [E,H,2I]`` (gate columns first), ``down_proj [E,I,H]``."""
    if x.shape[0] == 0:
        return x
    _C.require_cuda(x, gate_up_proj, down_proj, expert_ids)
    blk = _cached_block(gate_up_proj, down_proj, num_experts)
    out = blk.forward_routed(x, expert_ids.view(1, -1), [1.0], 1.0, 1.0, False,
                             torch.bfloat16 if x.dtype == torch.bfloat16 else torch.float32)
    return out if out.dtype == x.dtype else out.to(x.dtype)


def fused_token_routed_forward_int8(x: torch.Tensor, gate_up_int8: torch.Tensor, gate_up_scale: torch.Tensor,
                                    down_int8: torch.Tensor, down_scale: torch.Tensor, expert_ids: torch.Tensor,
                                    num_experts: int, intermediate_per_tp: int, bmm_threshold: int = 64
                                    ) -> torch.Tensor:
    """Weight-only INT8 with per-K-row scales (``gate_up_scale [E,H]``, ``down_scale [E,I]``,
    kernels/fused_experts.py:153-291, core/quantization.py:424-454): weights are dequantised once to the compute
    dtype (exactly what the reference's torch path does each call) and cached, then the bf16 grouped path runs."""
    if x.shape[0] == 0:
        return x
    _C.require_cuda(x, gate_up_int8, down_int8, expert_ids)
    srcs = (gate_up_int8, gate_up_scale, down_int8, down_scale)
    key = _key(*srcs)
    blk = _lookup(key, srcs)
    if blk is None:
        gu = (gate_up_int8.float() * gate_up_scale.unsqueeze(-1).float())
        dn = (down_int8.float() * down_scale.unsqueeze(-1).float())
        e, h, two_i = gu.shape
        inter = two_i // 2
        blk = I64Block(h, e, inter, 0, 1, 1, [1.0], 1.0, 1.0, _C.MODE_BF16, x.device)
        blk.load_bf16(gu[:, :, :inter], gu[:, :, inter:], dn)
        _store(key, srcs, blk)
    out = blk.forward_routed(x, expert_ids.view(1, -1), [1.0], 1.0, 1.0, False,
                             torch.bfloat16 if x.dtype == torch.bfloat16 else torch.float32)
    return out if out.dtype == x.dtype else out.to(x.dtype)


def fused_token_routed_forward_int4(x: torch.Tensor, gate_up_int4: torch.Tensor, gate_up_scale: torch.Tensor,
                                    gate_up_zero: torch.Tensor, down_int4: torch.Tensor, down_scale: torch.Tensor,
                                    down_zero: torch.Tensor, expert_ids: torch.Tensor, num_experts: int,
                                    intermediate_per_tp: int, *args, **kwargs) -> torch.Tensor:
    """Group-128 INT4 experts (kernels/fused_experts.py:294-476 of the reference): NOT built for sm_100a.  The name exists so
    that a caller gets this message rather than an ImportError or, worse, a silent PyTorch stand-in."""
    raise NotImplementedError(
        "vllm_i64 (B200 build): fused_token_routed_forward_int4 has no sm_100a kernel; use the INT8 (weight-only or W8A8) or FP8 "
        "experts, or no quantisation")
