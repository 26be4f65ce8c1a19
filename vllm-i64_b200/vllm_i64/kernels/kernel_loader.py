"""``get_ops()`` — the op object the reference obtains from ``kernels/kernel_loader.py:114-123``.

The reference returns either a JIT-compiled ``TORCH_LIBRARY(vllm_i64)`` module (csrc/i64_ops_binding.cu:400-407)
or ``FallbackOps``.  Here there is exactly one implementation: the prebuilt sm_100a library behind the C ABI.
Every op validates like the reference's ``TORCH_CHECK``s and raises ``RuntimeError`` on failure; calling an op
with CPU tensors raises (this build has no CPU path).
"""
from __future__ import annotations

from typing import Tuple

import torch

from vllm_i64 import _C


class I64Ops:
    """Op surface of ``TORCH_LIBRARY(vllm_i64)``: route_tokens, silu_hadamard, scatter_by_expert,
    gather_by_expert, atomic_scatter, fused_route_scatter."""

    @staticmethod
    def route_tokens(token_ids: torch.Tensor, num_experts: int) -> torch.Tensor:
        """``token_id & (E-1)`` → int32 (csrc/i64_ops_binding.cu:32-65)."""
        _C.require_cuda(token_ids)
        if token_ids.dtype != torch.int64:
            raise RuntimeError("token_ids must be int64")
        ids = token_ids.contiguous().view(-1)
        out = torch.empty(ids.numel(), dtype=torch.int32, device=ids.device)
        _C.check(_C.lib().i64b_route_modulo(ids.data_ptr(), out.data_ptr(), ids.numel(), int(num_experts),
                                            _C.stream_ptr(ids.device)))
        return out.view(token_ids.shape)

    @staticmethod
    def silu_hadamard(gate_up: torch.Tensor, expert_inter: int) -> torch.Tensor:
        """``silu(gate_up[:, :I]) * gate_up[:, I:]`` (csrc/i64_ops_binding.cu:74-172)."""
        _C.require_cuda(gate_up)
        if gate_up.dim() != 2 or gate_up.shape[1] != 2 * expert_inter:
            raise RuntimeError("gate_up must be [N, 2*expert_inter]")
        gu = gate_up.contiguous()
        out = torch.empty(gu.shape[0], expert_inter, dtype=gu.dtype, device=gu.device)
        _C.check(_C.lib().i64b_silu_hadamard(gu.data_ptr(), out.data_ptr(), gu.shape[0], int(expert_inter),
                                             _C.dtype_code(gu), _C.stream_ptr(gu.device)))
        return out

    @staticmethod
    def dispatch_plan(expert_ids: torch.Tensor, num_experts: int):
        """Deterministic plan for ``[K, N]`` (or ``[N]``) int32 expert ids: (scatter_indices, counts, offsets,
        dest_row, group_rows); the permutation equals a stable argsort per route."""
        _C.require_cuda(expert_ids)
        ids = expert_ids.to(torch.int32).contiguous()
        squeeze = ids.dim() == 1
        ids2 = ids.view(1, -1) if squeeze else ids
        k, n = ids2.shape
        dev = ids.device
        scatter = torch.empty(k, n, dtype=torch.int32, device=dev)
        dest = torch.empty(k, n, dtype=torch.int32, device=dev)
        counts = torch.empty(k, num_experts, dtype=torch.int32, device=dev)
        offsets = torch.empty(k, num_experts, dtype=torch.int32, device=dev)
        group_rows = torch.empty(num_experts, dtype=torch.int32, device=dev)
        lib = _C.lib()
        nbytes = lib.i64b_plan_workspace_bytes(n, k, int(num_experts))
        ws = _C.workspace(dev, nbytes)
        _C.check(lib.i64b_dispatch_plan(ids2.data_ptr(), k, n, int(num_experts), scatter.data_ptr(), counts.data_ptr(),
                                        offsets.data_ptr(), dest.data_ptr(), group_rows.data_ptr(), ws.data_ptr(),
                                        ws.numel(), _C.stream_ptr(dev)))
        if squeeze:
            return scatter[0], counts[0], offsets[0], dest[0], group_rows
        return scatter, counts, offsets, dest, group_rows

    @staticmethod
    def scatter_by_expert(hidden: torch.Tensor, expert_ids: torch.Tensor, num_experts: int
                          ) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor, torch.Tensor]:
        """(scattered, scatter_indices i32[N], expert_offsets i32[E], expert_counts i32[E]); order = stable sort
        by expert id (csrc/i64_ops_binding.cu:179-204)."""
        _C.require_cuda(hidden, expert_ids)
        if hidden.dim() != 2 or expert_ids.numel() != hidden.shape[0]:
            raise RuntimeError("hidden must be [N, H] and expert_ids [N]")
        h = hidden.contiguous()
        row_bytes = h.shape[1] * h.element_size()
        if row_bytes % 16 != 0:
            raise RuntimeError("hidden row size must be a multiple of 16 bytes")
        scatter, counts, offsets, _, _ = I64Ops.dispatch_plan(expert_ids.view(-1), num_experts)
        out = torch.empty_like(h)
        _C.check(_C.lib().i64b_scatter_rows(h.data_ptr(), scatter.data_ptr(), out.data_ptr(), h.shape[0], row_bytes,
                                            _C.stream_ptr(h.device)))
        return out, scatter, offsets, counts

    @staticmethod
    def gather_by_expert(expert_out: torch.Tensor, scatter_indices: torch.Tensor) -> torch.Tensor:
        """``expert_out[scatter_indices]`` (csrc/i64_ops_binding.cu:207-209)."""
        _C.require_cuda(expert_out, scatter_indices)
        src = expert_out.contiguous()
        idx = scatter_indices.to(torch.int32).contiguous()
        row_bytes = src.shape[1] * src.element_size()
        if row_bytes % 16 != 0:
            raise RuntimeError("row size must be a multiple of 16 bytes")
        out = torch.empty(idx.numel(), src.shape[1], dtype=src.dtype, device=src.device)
        _C.check(_C.lib().i64b_gather_rows(src.data_ptr(), idx.data_ptr(), out.data_ptr(), idx.numel(), row_bytes,
                                           _C.stream_ptr(src.device)))
        return out

    @staticmethod
    def atomic_scatter(hidden, expert_ids, num_experts):
        """Same 4-tuple as scatter_by_expert.  The reference claims slots with atomicAdd (order varies run to run,
        fp32 only — csrc/i64_ops_binding.cu:220-292); this build is deterministic and dtype-agnostic."""
        return I64Ops.scatter_by_expert(hidden, expert_ids, num_experts)

    @staticmethod
    def fused_route_scatter(token_ids, hidden, num_experts):
        """(expert_ids, scattered, scatter_map, offsets, counts) — csrc/i64_ops_binding.cu:302-393."""
        expert_ids = I64Ops.route_tokens(token_ids, num_experts)
        scattered, scatter_map, offsets, counts = I64Ops.scatter_by_expert(hidden, expert_ids, num_experts)
        return expert_ids, scattered, scatter_map, offsets, counts


_OPS = None


def get_ops() -> I64Ops:
    """The one op implementation of this build (raises ImportError if libi64b.so has not been built)."""
    global _OPS
    if _OPS is None:
        _C.lib()
        _OPS = I64Ops()
    return _OPS


def is_cuda_kernels_available() -> bool:
    try:
        _C.lib()
    except (ImportError, OSError):
        return False
    return torch.cuda.is_available()
