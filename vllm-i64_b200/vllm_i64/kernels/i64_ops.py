"""``vllm_i64.kernels.i64_ops`` — the functional surface of kernels/i64_ops.py:17-193 on the sm_100a kernels."""
from __future__ import annotations

from typing import Tuple

import torch

from vllm_i64.kernels.kernel_loader import get_ops


def i64_route_tokens(token_ids: torch.Tensor, num_experts: int) -> torch.Tensor:
    """``(token_ids & (E-1)).int32`` (kernels/i64_ops.py:17-34)."""
    return get_ops().route_tokens(token_ids, num_experts)


def i64_scatter(hidden_states: torch.Tensor, expert_ids: torch.Tensor, num_experts: int
                ) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor, torch.Tensor]:
    """(scattered, scatter_indices, expert_offsets, expert_counts) — kernels/i64_ops.py:37-78."""
    return get_ops().scatter_by_expert(hidden_states, expert_ids, num_experts)


def i64_gather(expert_output: torch.Tensor, scatter_indices: torch.Tensor) -> torch.Tensor:
    """kernels/i64_ops.py:81-97."""
    return get_ops().gather_by_expert(expert_output, scatter_indices)


def i64_expert_forward(scattered: torch.Tensor, gate_up_weights: torch.Tensor, down_weights: torch.Tensor,
                       expert_offsets: torch.Tensor, expert_counts: torch.Tensor, num_experts: int) -> torch.Tensor:
    """Per-segment ``silu(x@gate)·(x@up) @ down`` on expert-contiguous rows (kernels/i64_ops.py:100-158,
    csrc/i64_expert_dispatch.cu:144-233) — one grouped tcgen05 launch per GEMM, no host sync."""
    from vllm_i64.kernels.fused_experts import fused_token_routed_forward
    n = scattered.shape[0]
    ids = torch.repeat_interleave(torch.arange(num_experts, device=scattered.device),
                                  expert_counts.to(torch.int64), output_size=n)
    return fused_token_routed_forward(scattered, gate_up_weights, down_weights, ids, num_experts,
                                      down_weights.shape[1])


def i64_full_pipeline(token_ids: torch.Tensor, hidden_states: torch.Tensor, gate_up_weights: torch.Tensor,
                      down_weights: torch.Tensor, num_experts: int) -> torch.Tensor:
    """route → dispatch → experts → un-permute (kernels/i64_ops.py:161-193)."""
    from vllm_i64.kernels.fused_experts import fused_token_routed_forward
    expert_ids = i64_route_tokens(token_ids, num_experts)
    return fused_token_routed_forward(hidden_states, gate_up_weights, down_weights, expert_ids, num_experts,
                                      down_weights.shape[1])
