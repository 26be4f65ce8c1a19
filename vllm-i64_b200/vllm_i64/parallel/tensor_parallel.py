"""Tensor-parallel state and helpers the MoE / dense FFN layers use.

Mirror of the reference surface ``vllm_i64.parallel.tensor_parallel`` restricted to what the hot path touches:
``TPState``/``get_tp`` (parallel/tensor_parallel.py:33-42,75), ``init_distributed`` (:44-72),
``Column/RowParallelLinear`` (:91-160), ``shard_expert_weights`` (:164-192), ``all_reduce`` (:195-200).
"""
from __future__ import annotations

import os
from dataclasses import dataclass
from typing import Optional

import torch
import torch.distributed as dist
import torch.nn as nn


@dataclass
class TPState:
    tp_size: int = 1
    tp_rank: int = 0
    tp_group: Optional[dist.ProcessGroup] = None
    device: str = "cuda:0"


_TP = TPState()


def get_tp() -> TPState:
    return _TP


def get_tp_rank() -> int:
    return _TP.tp_rank


def get_tp_world_size() -> int:
    return _TP.tp_size


def init_distributed(tp_size: int = 1, backend: str = "nccl") -> None:
    """RANK / LOCAL_RANK from the environment (torchrun); one TP group per pipeline stage."""
    global _TP
    if tp_size <= 1:
        _TP = TPState(tp_size=1, tp_rank=0, device="cuda:0" if torch.cuda.is_available() else "cpu")
        return
    if not dist.is_initialized():
        dist.init_process_group(backend=backend)
    rank = dist.get_rank()
    local_rank = int(os.environ.get("LOCAL_RANK", rank))
    device = f"cuda:{local_rank}" if torch.cuda.is_available() else "cpu"
    if torch.cuda.is_available():
        torch.cuda.set_device(device)
    world = dist.get_world_size()
    stage = rank // tp_size
    group = None
    for s in range((world + tp_size - 1) // tp_size):   # every rank must create every group
        ranks = list(range(s * tp_size, min((s + 1) * tp_size, world)))
        g = dist.new_group(ranks)
        if s == stage:
            group = g
    _TP = TPState(tp_size=tp_size, tp_rank=rank % tp_size, tp_group=group, device=device)


def set_tp_state(state: TPState) -> None:
    global _TP
    _TP = state


class ColumnParallelLinear(nn.Module):
    """Output dim sharded across TP ranks (gate / up projections)."""

    def __init__(self, in_features: int, out_features: int, bias: bool = False):
        super().__init__()
        tp = get_tp()
        self.tp_size = tp.tp_size
        if out_features % tp.tp_size != 0:
            raise ValueError(f"out_features ({out_features}) not divisible by tp_size ({tp.tp_size})")
        self.out_per_rank = out_features // tp.tp_size
        self.linear = nn.Linear(in_features, self.out_per_rank, bias=bias)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return self.linear(x)

    def load_full_weight(self, full_weight: torch.Tensor, full_bias: Optional[torch.Tensor] = None) -> None:
        tp = get_tp()
        lo = tp.tp_rank * self.out_per_rank
        self.linear.weight.data.copy_(full_weight[lo:lo + self.out_per_rank])
        if full_bias is not None and self.linear.bias is not None:
            self.linear.bias.data.copy_(full_bias[lo:lo + self.out_per_rank])


class RowParallelLinear(nn.Module):
    """Input dim sharded across TP ranks; output all-reduced (down projection)."""

    def __init__(self, in_features: int, out_features: int, bias: bool = False):
        super().__init__()
        tp = get_tp()
        self.tp_size = tp.tp_size
        if in_features % tp.tp_size != 0:
            raise ValueError(f"in_features ({in_features}) not divisible by tp_size ({tp.tp_size})")
        self.in_per_rank = in_features // tp.tp_size
        self.linear = nn.Linear(self.in_per_rank, out_features, bias=False)
        self.bias = nn.Parameter(torch.zeros(out_features)) if bias else None

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        out = self.linear(x)
        if self.tp_size > 1:
            dist.all_reduce(out, group=get_tp().tp_group)
        if self.bias is not None:
            out = out + self.bias
        return out

    def load_full_weight(self, full_weight: torch.Tensor, full_bias: Optional[torch.Tensor] = None) -> None:
        tp = get_tp()
        lo = tp.tp_rank * self.in_per_rank
        self.linear.weight.data.copy_(full_weight[:, lo:lo + self.in_per_rank])
        if full_bias is not None and self.bias is not None:
            if tp.tp_rank == 0:
                self.bias.data.copy_(full_bias)
            else:
                self.bias.data.zero_()


def shard_expert_weights(gate_up: torch.Tensor, down: torch.Tensor):
    """Slice fused ``gate_up [E,H,2I]`` / ``down [E,I,H]`` on the intermediate dim for this TP rank."""
    tp = get_tp()
    if tp.tp_size <= 1:
        return gate_up, down
    full_inter = gate_up.shape[2] // 2
    if full_inter % tp.tp_size != 0:
        raise ValueError(
            f"Expert intermediate_size ({full_inter}) not divisible by tp_size ({tp.tp_size}). "
            f"Choose a tp_size that evenly divides {full_inter}.")
    per = full_inter // tp.tp_size
    lo = tp.tp_rank * per
    gate = gate_up[:, :, lo:lo + per]
    up = gate_up[:, :, full_inter + lo:full_inter + lo + per]
    return torch.cat([gate, up], dim=2).contiguous(), down[:, lo:lo + per, :].contiguous()


def all_reduce(tensor: torch.Tensor) -> torch.Tensor:
    """SUM over the TP group (no-op at tp_size 1)."""
    if _TP.tp_size <= 1:
        return tensor
    dist.all_reduce(tensor, group=_TP.tp_group)
    return tensor


from vllm_i64 import _overlay as _ov  # noqa: E402

_ov.fill(__name__, globals())
