"""Timed CPU baseline: a torch-CPU restatement of the reference's CPU serving path for the block.

TEST / BENCH INFRASTRUCTURE ONLY (see oracle/i64_oracle.py).  The reference's CPU recipe is
``TokenRoutedMLP`` (layers/token_routed_mlp.py:132-187, the per-expert mask loop) after
``_fuse_cpu_projection_layers`` (core/loader.py:79-94 → fuse_shared_gate_up / materialize_expert_linears) and
``_quantize_cpu_dynamic`` (core/loader.py:57-76: ``torch.ao.quantization.quantize_dynamic`` on every
``nn.Linear`` with the x86/FBGEMM engine), with threads set as ``vllm-i64 serve`` sets them (cli.py:116-134).
The reference itself is Python and cannot travel to the GPU box, so this port is what ``bench.py`` times there
(``cpu_baseline.kind = "port"``).  Parity of the dynamic-INT8 arithmetic is UNPINNED by design — it is a timing
baseline, not an oracle (FBGEMM's per-tensor asymmetric activations differ from fp32 by ~9 % mean-rel);
the fp32 mode of this module IS pinned (tests/test_oracle_golden.py::test_cpu_reference_port_matches_oracle).
"""
from __future__ import annotations

import os
from typing import Optional

import torch
import torch.nn as nn
import torch.nn.functional as F


class CpuTokenRoutedMLP(nn.Module):
    """Materialised-linear form of the block (what the reference's CPU engine runs)."""

    def __init__(self, hidden: int, num_experts: int, expert_inter: int, shared_inter: int, top_k: int,
                 primary_weight: float, routed_scale: float, shared_scale: float, table: torch.Tensor):
        super().__init__()
        self.hidden, self.num_experts, self.top_k = hidden, num_experts, top_k
        self.primary_weight = primary_weight if top_k > 1 else 1.0
        self.routed_scale, self.shared_scale = routed_scale, shared_scale
        self.expert_gate_up = nn.ModuleList(nn.Linear(hidden, 2 * expert_inter, bias=False) for _ in range(num_experts))
        self.expert_down = nn.ModuleList(nn.Linear(expert_inter, hidden, bias=False) for _ in range(num_experts))
        self.has_shared = shared_inter > 0
        if self.has_shared:
            self.shared_gate_up = nn.Linear(hidden, 2 * shared_inter, bias=False)
            self.shared_down = nn.Linear(shared_inter, hidden, bias=False)
        self.register_buffer("topk_token_to_expert", table.to(torch.long))

    @torch.no_grad()
    def load(self, gate, up, down, shared_gate=None, shared_up=None, shared_down=None):
        """gate/up ``[E,H,I]``, down ``[E,I,H]`` → (2I,H) / (H,I) linears (token_routed_mlp.py:232-242)."""
        for e in range(self.num_experts):
            self.expert_gate_up[e].weight.copy_(torch.cat([gate[e].T, up[e].T], dim=0))
            self.expert_down[e].weight.copy_(down[e].T)
        if self.has_shared:
            self.shared_gate_up.weight.copy_(torch.cat([shared_gate, shared_up], dim=0))
            self.shared_down.weight.copy_(shared_down)
        return self

    def expert_forward(self, x: torch.Tensor, ids: torch.Tensor) -> torch.Tensor:
        out = torch.zeros_like(x)
        for e in range(self.num_experts):
            mask = ids == e
            if not mask.any():
                continue
            g, u = self.expert_gate_up[e](x[mask]).chunk(2, dim=-1)
            out[mask] = self.expert_down[e](F.silu(g) * u).to(out.dtype)
        return out

    def forward(self, x: torch.Tensor, token_ids: torch.Tensor) -> torch.Tensor:
        v = self.topk_token_to_expert.shape[1]
        routes = self.topk_token_to_expert[:, token_ids.clamp(0, v - 1)]
        out = self.primary_weight * self.expert_forward(x, routes[0])
        if self.top_k > 1:
            sec = (1.0 - self.primary_weight) / (self.top_k - 1)
            for r in range(1, self.top_k):
                out = out + sec * self.expert_forward(x, routes[r])
        if self.has_shared:
            g, u = self.shared_gate_up(x).chunk(2, dim=-1)
            out = self.routed_scale * out + self.shared_scale * self.shared_down(F.silu(g) * u).to(out.dtype)
        else:
            out = self.routed_scale * out
        return out


def quantize_cpu_dynamic(module: nn.Module) -> nn.Module:
    """core/loader.py:57-76 — dynamic INT8 on nn.Linear with the x86 (or fbgemm) engine."""
    engines = torch.backends.quantized.supported_engines
    engine = "x86" if "x86" in engines else "fbgemm" if "fbgemm" in engines else None
    if engine is None:
        raise RuntimeError(f"CPU INT8 needs the x86 or FBGEMM quantized backend; available: {engines}")
    torch.backends.quantized.engine = engine
    module.eval()
    return torch.ao.quantization.quantize_dynamic(module, {nn.Linear}, dtype=torch.qint8, inplace=False)


def set_cpu_threads(requested: Optional[int] = None) -> int:
    """cli.py:116-134: min(VLLM_I64_CPU_THREADS, affinity), inter-op threads 1."""
    avail = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    want = requested or int(os.environ.get("VLLM_I64_CPU_THREADS", avail))
    n = max(1, min(want, avail))
    torch.set_num_threads(n)
    try:
        torch.set_num_interop_threads(1)
    except RuntimeError:
        pass
    return n
