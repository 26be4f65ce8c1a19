"""CPU oracle for the token-routed MoE FFN block of Complexity-ML/vllm-i64.

TEST INFRASTRUCTURE ONLY.  Nothing under ``vllm-i64_b200/`` (the product) may
import this module; only ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` do, and there only
as the checker or as the timed CPU baseline.

This is a *restatement* of the reference's algorithm (numpy for the
integer/byte arithmetic, CPU torch only for the BLAS-backed fp32 matmuls),
written against the behaviour of the reference files cited on every function
(paths relative to the reference checkout).  Parity status: PINNED — every
function here is checked against golden vectors produced by importing the
reference itself in the build container (``tests/golden/make_golden.py`` →
``tests/golden/*.npz``) and against the reference's own known-answer tests
(``tests/test_i64_pipeline.py:35-57``, ``tests/test_competitive_features.py:491-511``,
``tests/test_loader.py:219-246``, ``tests/test_model.py:198-205``); see
``tests/test_oracle_golden.py``.

Third-party arithmetic on the path: ``torch`` (unpinned ``>=2.0`` in the
reference's ``pyproject.toml:26``; 2.11.0 here).  ``torch._int_mm`` is exact
integer arithmetic (restated below through an exact float64 BLAS product);
``torch.ao.quantization.quantize_dynamic`` (FBGEMM/x86) is only the *timed*
CPU baseline of config 1 and is deliberately not a parity oracle (it differs
from fp32 by ~9 % mean-rel on random weights).
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

Q7 = 128  # layers/integer_activations.py:18
SILU_LUT_MIN = -1024  # layers/integer_activations.py:25
SILU_LUT_MAX = 1024  # layers/integer_activations.py:26


# --------------------------------------------------------------------------- #
# a1 / a6 — routing
# --------------------------------------------------------------------------- #

def route_table(token_ids: Optional[np.ndarray], table: np.ndarray, num_tokens: int,
                num_experts: int) -> np.ndarray:
    """``TokenRoutedMLP.route`` (layers/token_routed_mlp.py:113-130).

    ``table`` is ``topk_token_to_expert`` ``[K, V]``.  Ids are clamped to
    ``[0, V-1]`` (:123) and gathered column-wise; ``token_ids is None`` routes
    every token to experts ``(0 + k) % E`` (:120-130).  Returns int64 ``[K, N]``.
    """
    table = np.asarray(table, dtype=np.int64)
    k, v = table.shape
    if token_ids is None:
        return np.stack([np.full(num_tokens, r % num_experts, dtype=np.int64) for r in range(k)])
    ids = np.clip(np.asarray(token_ids, dtype=np.int64), 0, v - 1)
    return table[:, ids]


def default_tables(vocab_size: int, num_experts: int, top_k: int) -> Tuple[np.ndarray, np.ndarray]:
    """Constructor-time buffers (layers/token_routed_mlp.py:92-107):
    ``token_to_expert = arange(V) % E`` and the cyclic top-k stack."""
    ar = np.arange(vocab_size, dtype=np.int64)
    return ar % num_experts, np.stack([(ar + r) % num_experts for r in range(top_k)])


def legacy_table_to_topk(token_to_expert: np.ndarray, num_experts: int, top_k: int) -> np.ndarray:
    """Loader fallback for checkpoints that only ship the primary route
    (core/loader.py:389-407): route r is ``(primary + r) % E``."""
    p = np.asarray(token_to_expert, dtype=np.int64)
    return np.stack([(p + r) % num_experts for r in range(top_k)])


def route_modulo(token_ids: np.ndarray, num_experts: int) -> np.ndarray:
    """``i64_route_tokens`` (kernels/i64_ops.py:17-34; csrc/i64_router.cu:51-70):
    ``token_id & (E-1)`` as int32."""
    return (np.asarray(token_ids, dtype=np.int64) & (num_experts - 1)).astype(np.int32)


def synthetic_hash_table(vocab_size: int, num_experts: int, top_k: int, layer: int) -> np.ndarray:
    """Synthetic layer-specific balanced hash table (not a reference function): see tools/synth_data.py."""
    import os
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    if root not in sys.path:
        sys.path.insert(0, root)
    from tools.synth_data import synthetic_hash_table as gen
    return gen(vocab_size, num_experts, top_k, layer)


# --------------------------------------------------------------------------- #
# a7 / a8 — dispatch plan (stable sort by expert) and gather
# --------------------------------------------------------------------------- #

@dataclass
class DispatchPlan:
    perm: np.ndarray             # [N] token index stored at each sorted row
    scatter_indices: np.ndarray  # [N] sorted row of each token
    counts: np.ndarray           # [E]
    offsets: np.ndarray          # [E] exclusive scan
    rank: np.ndarray             # [N] stable rank of each token among its expert's tokens


def dispatch_plan(expert_ids: np.ndarray, num_experts: int) -> DispatchPlan:
    """Canonical dispatch order = stable sort by expert id
    (csrc/i64_ops_binding.cu:179-204 ``scatter_by_expert``;
    kernels/fused_experts.py:112-119 ``argsort(stable=True)`` → bincount →
    cumsum; kernels/i64_ops.py:37-78).  ``scatter_indices[t]`` is the row token
    ``t`` lands in."""
    ids = np.asarray(expert_ids).astype(np.int64)
    n = ids.shape[0]
    perm = np.argsort(ids, kind="stable").astype(np.int64)
    counts = np.bincount(ids, minlength=num_experts).astype(np.int64)[:num_experts]
    offsets = np.zeros(num_experts, dtype=np.int64)
    if num_experts > 1:
        offsets[1:] = np.cumsum(counts)[:-1]
    scatter = np.empty(n, dtype=np.int64)
    scatter[perm] = np.arange(n, dtype=np.int64)
    rank = scatter - offsets[ids] if n else scatter
    return DispatchPlan(perm, scatter, counts, offsets, rank)


def scatter_rows(hidden: np.ndarray, plan: DispatchPlan) -> np.ndarray:
    """``scattered = hidden[sorted_indices]`` (kernels/i64_ops.py:69-70)."""
    return hidden[plan.perm]


def gather_rows(expert_out: np.ndarray, scatter_indices: np.ndarray) -> np.ndarray:
    """``i64_gather`` (kernels/i64_ops.py:81-97; csrc/i64_router.cu:150-165)."""
    return expert_out[np.asarray(scatter_indices, dtype=np.int64)]


@dataclass
class MergedLayout:
    """Expert-major layout used by the B200 grouped GEMM: for expert e the rows
    of route 0 come first, then route 1, ... each in stable token order.  A pure
    function of the per-route plans above (so bit-exactness of the plans implies
    bit-exactness of this layout)."""
    group_start: np.ndarray   # [E]
    group_rows: np.ndarray    # [E]
    dest_row: np.ndarray      # [K, N]


def merged_layout(routes: np.ndarray, num_experts: int) -> Tuple[MergedLayout, List[DispatchPlan]]:
    routes = np.asarray(routes, dtype=np.int64)
    k, n = routes.shape
    plans = [dispatch_plan(routes[r], num_experts) for r in range(k)]
    counts = np.stack([p.counts for p in plans])                     # [K, E]
    group_rows = counts.sum(axis=0)
    group_start = np.zeros(num_experts, dtype=np.int64)
    group_start[1:] = np.cumsum(group_rows)[:-1]
    route_base = np.zeros_like(counts)
    route_base[1:] = np.cumsum(counts, axis=0)[:-1]
    dest = np.empty((k, n), dtype=np.int64)
    for r in range(k):
        e = routes[r]
        dest[r] = group_start[e] + route_base[r, e] + plans[r].rank
    return MergedLayout(group_start, group_rows, dest), plans


# --------------------------------------------------------------------------- #
# a12 — W8A8 arithmetic
# --------------------------------------------------------------------------- #

def quantize_rows_int8(x: np.ndarray) -> Tuple[np.ndarray, np.ndarray]:
    """Per-row symmetric INT8: ``quantize_activations_int8``
    (core/quantization.py:103-116), ``quantize_int8`` on fp32 weights (:80-91),
    csrc/I64_quantize.cu:41-92.  All arithmetic in IEEE fp32:
    ``s = max(absmax, 1e-8)/127``; ``q = clamp(round_half_even(x / s), -128, 127)``."""
    x32 = np.ascontiguousarray(x, dtype=np.float32)
    absmax = np.maximum(np.max(np.abs(x32), axis=-1, keepdims=True), np.float32(1e-8)).astype(np.float32)
    scale = (absmax / np.float32(127.0)).astype(np.float32)
    q = np.clip(np.rint((x32 / scale).astype(np.float32)), -128, 127).astype(np.int8)
    return q, scale[..., 0].copy()


def dequantize_int8(q: np.ndarray, scale: np.ndarray) -> np.ndarray:
    """core/quantization.py:94-96; csrc/I64_quantize.cu:99-111."""
    return q.astype(np.float32) * np.asarray(scale, dtype=np.float32)[..., None]


def int_mm(a_i8: np.ndarray, bt_i8: np.ndarray) -> np.ndarray:
    """``torch._int_mm(a, b)`` with ``b = bt.T`` (core/quantization.py:168-172):
    exact int8×int8→int32.  Evaluated through a float64 BLAS product, which is
    exact here because |acc| ≤ K·128·128 < 2**53."""
    acc = a_i8.astype(np.float64) @ bt_i8.astype(np.float64).T
    return acc.astype(np.int64).astype(np.int32)


def rescale_acc(acc_i32: np.ndarray, x_scale: np.ndarray, w_scale: np.ndarray) -> np.ndarray:
    """``result_i32.float() * (x_scale[:, None] * w_scale[None, :])``
    (core/quantization.py:173,278) — the scale product is formed first."""
    s = (np.asarray(x_scale, np.float32)[:, None] * np.asarray(w_scale, np.float32)[None, :]).astype(np.float32)
    return (acc_i32.astype(np.float32) * s).astype(np.float32)


_SILU_LUT_CACHE: Optional[np.ndarray] = None


def silu_lut_q7() -> np.ndarray:
    """``_build_silu_lut`` (layers/integer_activations.py:28-34):
    ``round(silu(i/128)*128)`` for i in [-1024, 1024], int32.  Built with the same
    torch fp32 ops as the reference so the table is bit-identical (pinned by
    tests/golden/silu_lut.npy)."""
    global _SILU_LUT_CACHE
    if _SILU_LUT_CACHE is None:
        import torch
        import torch.nn.functional as F
        idx = torch.arange(SILU_LUT_MIN, SILU_LUT_MAX + 1, dtype=torch.float32)
        _SILU_LUT_CACHE = (F.silu(idx / Q7) * Q7).round().to(torch.int32).numpy().copy()
    return _SILU_LUT_CACHE


def silu_q7(x_q7: np.ndarray) -> np.ndarray:
    """``silu_integer`` (layers/integer_activations.py:39-51)."""
    lut = silu_lut_q7()
    x = np.asarray(x_q7, dtype=np.int32)
    r = lut[(np.clip(x, SILU_LUT_MIN, SILU_LUT_MAX) - SILU_LUT_MIN).astype(np.int64)]
    r = np.where(x > SILU_LUT_MAX, x, r)
    r = np.where(x < SILU_LUT_MIN, 0, r)
    return r.astype(np.int32)


def silu_mul_q14(gate: np.ndarray, up: np.ndarray) -> np.ndarray:
    """Integer part of ``silu_multiply_integer`` (layers/integer_activations.py:53-66):
    Q7(gate) → LUT → × Q7(up) → Q14 int32."""
    g = np.rint(np.asarray(gate, np.float32) * np.float32(Q7)).astype(np.int32)
    u = np.rint(np.asarray(up, np.float32) * np.float32(Q7)).astype(np.int32)
    return (silu_q7(g) * u).astype(np.int32)


def silu_mul_integer(gate: np.ndarray, up: np.ndarray) -> np.ndarray:
    """``silu_multiply_integer`` → fp32 (``inter_q14.float() / 128**2``)."""
    return (silu_mul_q14(gate, up).astype(np.float32) / np.float32(Q7 * Q7)).astype(np.float32)


@dataclass
class W8A8Trace:
    x_q: np.ndarray
    x_scale: np.ndarray
    acc1: np.ndarray
    inter_q14: np.ndarray
    inter: np.ndarray
    inter_q: np.ndarray
    inter_scale: np.ndarray
    acc2: np.ndarray
    y: np.ndarray


def w8a8_swiglu(x: np.ndarray, w13_q: np.ndarray, w13_s: np.ndarray, w2_q: np.ndarray, w2_s: np.ndarray,
                x_preq: Optional[Tuple[np.ndarray, np.ndarray]] = None) -> W8A8Trace:
    """One SwiGLU expert in dynamic W8A8: ``MoEExpert._forward_int8_fused``
    (layers/moe.py:114-122) == ``DenseMLP._forward_int8_fused``
    (layers/dense_mlp.py:85-97):  act-quant → fused gate|up ``_int_mm`` → rescale →
    LUT-SiLU·up → re-quant → down ``_int_mm`` → rescale.
    ``w13_q`` is ``(2I, H)`` int8 = cat(gate, up) rows; ``w2_q`` is ``(H, I)``."""
    if x_preq is None:
        x_q, x_s = quantize_rows_int8(x)
    else:
        x_q, x_s = x_preq
    inter_size = w13_q.shape[0] // 2
    acc1 = int_mm(x_q, w13_q)
    gu = rescale_acc(acc1, x_s, w13_s)
    q14 = silu_mul_q14(gu[:, :inter_size], gu[:, inter_size:])
    inter = (q14.astype(np.float32) / np.float32(Q7 * Q7)).astype(np.float32)
    i_q, i_s = quantize_rows_int8(inter)
    acc2 = int_mm(i_q, w2_q)
    y = rescale_acc(acc2, i_s, w2_s)
    return W8A8Trace(x_q, x_s, acc1, q14, inter, i_q, i_s, acc2, y)


# --------------------------------------------------------------------------- #
# a2 / a3 / a5 — the block in floating point
# --------------------------------------------------------------------------- #

def _silu(x: np.ndarray) -> np.ndarray:
    x = x.astype(np.float32)
    return (x / (np.float32(1.0) + np.exp(-x, dtype=np.float32))).astype(np.float32)


def _mm(a: np.ndarray, b: np.ndarray) -> np.ndarray:
    return np.matmul(a.astype(np.float32), b.astype(np.float32))


def round_bf16(a: np.ndarray) -> np.ndarray:
    """fp32 → nearest bf16 (ties to even) → fp32, in pure numpy."""
    u = np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)
    r = ((u + np.uint32(0x7FFF) + ((u >> np.uint32(16)) & np.uint32(1))) & np.uint32(0xFFFF0000)).astype(np.uint32)
    return r.view(np.float32)


def _inter(x: np.ndarray, inter_bf16: bool) -> np.ndarray:
    """The reference's bf16 path materialises the SwiGLU intermediate as a bf16 tensor
    (layers/token_routed_mlp.py:150 with bf16 operands); ``inter_bf16=True`` reproduces that single rounding
    so the comparison with a bf16 implementation is not dominated by it."""
    return round_bf16(x) if inter_bf16 else x


@dataclass
class MoEWeights:
    """Tensors exactly as the reference module holds them
    (layers/token_routed_mlp.py:66-74,80-82): ``gate/up [E,H,I]``, ``down [E,I,H]``,
    shared ``nn.Linear`` weights ``(out, in)``."""
    gate: np.ndarray
    up: np.ndarray
    down: np.ndarray
    shared_gate: Optional[np.ndarray] = None
    shared_up: Optional[np.ndarray] = None
    shared_down: Optional[np.ndarray] = None


@dataclass
class MoEConfig:
    hidden_size: int
    num_experts: int
    expert_inter: int
    vocab_size: int
    top_k: int = 1
    primary_weight: float = 1.0
    shared_inter: int = 0
    shared_output_scale: float = 1.0
    routed_output_scale: float = 1.0

    def route_weights(self) -> List[float]:
        """layers/token_routed_mlp.py:53-58,156-158: w0 = primary, others share (1-w0)."""
        if self.top_k <= 1:
            return [1.0]
        sec = (1.0 - self.primary_weight) / (self.top_k - 1)
        return [self.primary_weight] + [sec] * (self.top_k - 1)


def expert_forward_fp32(x: np.ndarray, ids: np.ndarray, w: MoEWeights, num_experts: int,
                        inter_bf16: bool = False) -> np.ndarray:
    """``TokenRoutedMLP.expert_forward`` (layers/token_routed_mlp.py:132-152), tp=1."""
    out = np.zeros_like(x, dtype=np.float32)
    for e in range(num_experts):
        rows = np.nonzero(ids == e)[0]
        if rows.size == 0:
            continue
        xe = x[rows].astype(np.float32)
        inter = _inter(_silu(_mm(xe, w.gate[e])) * _mm(xe, w.up[e]), inter_bf16)
        out[rows] = _mm(inter, w.down[e])
    return out


def shared_forward_fp32(x: np.ndarray, w: MoEWeights, inter_bf16: bool = False) -> np.ndarray:
    """Shared SwiGLU (layers/token_routed_mlp.py:165-173); Linear weights are (out, in)."""
    x = x.astype(np.float32)
    inter = _inter(_silu(_mm(x, w.shared_gate.T)) * _mm(x, w.shared_up.T), inter_bf16)
    return _mm(inter, w.shared_down.T)


def combine_fp32(routed: Sequence[np.ndarray], shared: Optional[np.ndarray], cfg: MoEConfig) -> np.ndarray:
    """Top-k weighted sum + output scales in the reference's fp32 operation order
    (layers/token_routed_mlp.py:156-185): ``o = w0*y0; o = o + wk*yk; o = rs*o + ss*shared``."""
    ws = cfg.route_weights()
    o = (np.float32(ws[0]) * routed[0].astype(np.float32)).astype(np.float32)
    for r in range(1, len(routed)):
        o = (o + (np.float32(ws[r]) * routed[r].astype(np.float32)).astype(np.float32)).astype(np.float32)
    o = (np.float32(cfg.routed_output_scale) * o).astype(np.float32)
    if shared is not None:
        o = (o + (np.float32(cfg.shared_output_scale) * shared.astype(np.float32)).astype(np.float32)).astype(np.float32)
    return o


def moe_block_fp32(x: np.ndarray, token_ids: Optional[np.ndarray], table: np.ndarray, w: MoEWeights,
                   cfg: MoEConfig, inter_bf16: bool = False, routed_bf16: bool = False) -> np.ndarray:
    """``TokenRoutedMLP.forward`` (layers/token_routed_mlp.py:154-187) in fp32.

    ``routed_bf16``: the expert-parallel form of the same block.  The reference's EP path sends every routed expert's output
    back to the token's home rank as a tensor of the MODEL dtype (layers/token_routed_i64/layer.py:283-292: ``result_sorted``
    is allocated with ``dtype=dtype`` and filled by ``all_to_all_single``), so with a bf16 model each per-assignment output
    row is rounded to bf16 once before the home rank weighs and sums it; the shared expert never travels and is not
    rounded.  This is the oracle for the bf16-transport EP mode."""
    routes = route_table(token_ids, table, x.shape[0], cfg.num_experts)
    routed = [expert_forward_fp32(x, routes[r], w, cfg.num_experts, inter_bf16) for r in range(cfg.top_k)]
    if routed_bf16:
        routed = [round_bf16(r) for r in routed]
    shared = shared_forward_fp32(x, w, inter_bf16) if w.shared_gate is not None else None
    return combine_fp32(routed, shared, cfg)


def dense_swiglu_fp32(x: np.ndarray, gate_w: np.ndarray, up_w: np.ndarray, down_w: np.ndarray,
                      inter_bf16: bool = False) -> np.ndarray:
    """``DenseSwiGLUMLP.forward`` (models/complexity_deep/model.py:250-255) /
    ``DenseMLP.forward`` float branch (layers/dense_mlp.py:59-73); weights (out, in)."""
    x = x.astype(np.float32)
    return _mm(_inter(_silu(_mm(x, gate_w.T)) * _mm(x, up_w.T), inter_bf16), down_w.T)


def expert_segments_fp32(scattered: np.ndarray, gate_up: np.ndarray, down: np.ndarray,
                         offsets: np.ndarray, counts: np.ndarray, inter_bf16: bool = False) -> np.ndarray:
    """``i64_expert_forward`` (kernels/i64_ops.py:100-158; csrc/i64_expert_dispatch.cu:144-233):
    fused ``gate_up [E,H,2I]`` (gate columns first), ``down [E,I,H]``."""
    out = np.zeros((scattered.shape[0], down.shape[2]), dtype=np.float32)
    inter = down.shape[1]
    for e in range(gate_up.shape[0]):
        c, o = int(counts[e]), int(offsets[e])
        if c == 0:
            continue
        gu = _mm(scattered[o:o + c], gate_up[e])
        out[o:o + c] = _mm(_inter(_silu(gu[:, :inter]) * gu[:, inter:], inter_bf16), down[e])
    return out


def fused_token_routed_fp32(x: np.ndarray, gate_up: np.ndarray, down: np.ndarray, expert_ids: np.ndarray,
                            num_experts: int, inter_bf16: bool = False) -> np.ndarray:
    """``fused_token_routed_forward`` (kernels/fused_experts.py:29-150): sort → per
    expert SwiGLU → unsort.  Both of the reference's modes compute this."""
    plan = dispatch_plan(expert_ids, num_experts)
    seg = expert_segments_fp32(x[plan.perm], gate_up, down, plan.offsets, plan.counts, inter_bf16)
    return seg[plan.scatter_indices]


# --------------------------------------------------------------------------- #
# W8A8 block (composite oracle)
# --------------------------------------------------------------------------- #

@dataclass
class MoEWeightsInt8:
    """Per expert ``w13 (2I,H)`` = cat(gate[e].T, up[e].T) and ``w2 (H,I)`` = down[e].T
    (layers/token_routed_mlp.py:232-242) quantised per out-row with ``quantize_int8``."""
    w13_q: np.ndarray   # [E, 2I, H] int8
    w13_s: np.ndarray   # [E, 2I]
    w2_q: np.ndarray    # [E, H, I] int8
    w2_s: np.ndarray    # [E, H]
    sh13_q: Optional[np.ndarray] = None  # [2S, H]
    sh13_s: Optional[np.ndarray] = None
    sh2_q: Optional[np.ndarray] = None   # [H, S]
    sh2_s: Optional[np.ndarray] = None


def quantize_moe_weights(w: MoEWeights) -> MoEWeightsInt8:
    e = w.gate.shape[0]
    w13 = np.concatenate([np.transpose(w.gate, (0, 2, 1)), np.transpose(w.up, (0, 2, 1))], axis=1)
    w2 = np.transpose(w.down, (0, 2, 1))
    q13, s13, q2, s2 = [], [], [], []
    for i in range(e):
        q, s = quantize_rows_int8(w13[i]); q13.append(q); s13.append(s)
        q, s = quantize_rows_int8(w2[i]); q2.append(q); s2.append(s)
    out = MoEWeightsInt8(np.stack(q13), np.stack(s13), np.stack(q2), np.stack(s2))
    if w.shared_gate is not None:
        out.sh13_q, out.sh13_s = quantize_rows_int8(np.concatenate([w.shared_gate, w.shared_up], axis=0))
        out.sh2_q, out.sh2_s = quantize_rows_int8(w.shared_down)
    return out


def moe_block_w8a8(x: np.ndarray, token_ids: Optional[np.ndarray], table: np.ndarray, wq: MoEWeightsInt8,
                   cfg: MoEConfig, trace: Optional[Dict] = None) -> np.ndarray:
    """a3 with every expert (routed and shared) evaluated by :func:`w8a8_swiglu`.
    ``trace`` (optional dict) receives the bit-exact intermediates laid out in
    the merged expert-major order (x_q, x_scale, acc1, inter_q14, inter_q,
    inter_scale, acc2 per dispatched row)."""
    n = x.shape[0]
    routes = route_table(token_ids, table, n, cfg.num_experts)
    x_q, x_s = quantize_rows_int8(x)
    layout, _ = merged_layout(routes, cfg.num_experts)
    total = n * cfg.top_k
    h = cfg.hidden_size
    i_e = wq.w2_q.shape[2]
    if trace is not None:
        trace.update(x_q=np.zeros((total, h), np.int8), x_scale=np.zeros(total, np.float32),
                     acc1=np.zeros((total, 2 * i_e), np.int32), inter_q14=np.zeros((total, i_e), np.int32),
                     inter_q=np.zeros((total, i_e), np.int8), inter_scale=np.zeros(total, np.float32),
                     acc2=np.zeros((total, h), np.int32), y=np.zeros((total, h), np.float32),
                     dest_row=layout.dest_row.copy(), routes=routes.copy(),
                     tok_x_q=x_q.copy(), tok_x_scale=x_s.copy())
    routed = []
    for r in range(cfg.top_k):
        y = np.zeros((n, h), dtype=np.float32)
        for e in range(cfg.num_experts):
            rows = np.nonzero(routes[r] == e)[0]
            if rows.size == 0:
                continue
            t = w8a8_swiglu(None, wq.w13_q[e], wq.w13_s[e], wq.w2_q[e], wq.w2_s[e],
                            x_preq=(x_q[rows], x_s[rows]))
            y[rows] = t.y
            if trace is not None:
                d = layout.dest_row[r, rows]
                trace["x_q"][d] = t.x_q; trace["x_scale"][d] = t.x_scale; trace["acc1"][d] = t.acc1
                trace["inter_q14"][d] = t.inter_q14; trace["inter_q"][d] = t.inter_q
                trace["inter_scale"][d] = t.inter_scale; trace["acc2"][d] = t.acc2; trace["y"][d] = t.y
        routed.append(y)
    shared = None
    if wq.sh13_q is not None:
        ts = w8a8_swiglu(None, wq.sh13_q, wq.sh13_s, wq.sh2_q, wq.sh2_s, x_preq=(x_q, x_s))
        shared = ts.y
        if trace is not None:
            trace.update(sh_acc1=ts.acc1, sh_inter_q14=ts.inter_q14, sh_inter_q=ts.inter_q,
                         sh_inter_scale=ts.inter_scale, sh_acc2=ts.acc2, sh_y=ts.y)
    return combine_fp32(routed, shared, cfg)


# --------------------------------------------------------------------------- #
# f4 — FP8 E4M3 block (core/fp8.py; layers/dense_mlp.py:75-83)
# --------------------------------------------------------------------------- #

FP8_MAX = np.float32(448.0)


def e4m3_round(x: np.ndarray) -> np.ndarray:
    """fp32 -> nearest ``float8_e4m3fn`` value (round to nearest even), returned as fp32.  4 exponent bits (bias 7),
    3 mantissa bits, no infinities: normals 2^-6 .. 448 with a quantum of 2^(e-3), subnormals in steps of 2^-9.
    Inputs are expected inside [-448, 448] (the reference clamps first, core/fp8.py:99-102,127)."""
    x = np.asarray(x, dtype=np.float32)
    _, ex = np.frexp(x)                                # x = m * 2^ex, 0.5 <= |m| < 1  ->  floor(log2|x|) = ex - 1
    e = np.maximum(ex.astype(np.int32) - 1, -6)         # below 2^-6 the subnormal quantum 2^-9 applies
    quantum = np.ldexp(np.float32(1.0), e - 3).astype(np.float32)
    q = (np.rint(x.astype(np.float64) / quantum) * quantum).astype(np.float32)
    return np.clip(q, -FP8_MAX, FP8_MAX)


def e4m3_encode(v: np.ndarray) -> np.ndarray:
    """Byte pattern (uint8) of values that are already exactly representable in e4m3fn (output of e4m3_round)."""
    v = np.asarray(v, dtype=np.float32)
    sign = (np.signbit(v)).astype(np.uint8) << 7
    a = np.abs(v)
    m, ex = np.frexp(a)
    e = ex.astype(np.int32) - 1
    normal = a >= np.float32(2.0 ** -6)
    mant_n = np.rint((a / np.ldexp(np.float32(1.0), e) - 1.0) * 8).astype(np.int32)
    mant_s = np.rint(a / np.float32(2.0 ** -9)).astype(np.int32)
    byte = np.where(normal, ((e + 7) << 3) | mant_n, mant_s).astype(np.uint8)
    return (sign | byte).astype(np.uint8)


def quantize_rows_fp8(x: np.ndarray) -> Tuple[np.ndarray, np.ndarray]:
    """``quantize_activations_fp8`` (core/fp8.py:113-129) and ``quantize_fp8(per_channel=True)`` (:75-110): per row
    ``scale = max(absmax, 1e-12) / 448`` and ``q = e4m3((x / scale).clamp(-448, 448))``, all in IEEE fp32.
    Returns ``(q as fp32 values, scale [rows, 1])``."""
    x32 = np.ascontiguousarray(x, dtype=np.float32)
    absmax = np.maximum(np.max(np.abs(x32), axis=-1, keepdims=True), np.float32(1e-12)).astype(np.float32)
    scale = (absmax / FP8_MAX).astype(np.float32)
    q = e4m3_round(np.clip((x32 / scale).astype(np.float32), -FP8_MAX, FP8_MAX))
    return q, scale


def scaled_mm(a_q: np.ndarray, a_s: np.ndarray, bt_q: np.ndarray, b_s: np.ndarray) -> np.ndarray:
    """``torch._scaled_mm(x_fp8, w_fp8.T, scale_a=[M,1], scale_b=[1,N])`` (core/fp8.py:176-182): the fp8 products are exact
    in fp32, the sum is the tensor core's (fp64 here: the comparison is by tolerance), then ``* (scale_a * scale_b)``."""
    acc = (a_q.astype(np.float64) @ bt_q.astype(np.float64).T).astype(np.float32)
    return (acc * (a_s.reshape(-1, 1).astype(np.float32) * b_s.reshape(1, -1).astype(np.float32))).astype(np.float32)


def fp8_swiglu(x: Optional[np.ndarray], w13_q, w13_s, w2_q, w2_s, x_preq=None) -> np.ndarray:
    """One expert in FP8 (layers/dense_mlp.py:75-83): ``fp8_fused_gate_up`` -> ``silu(gate) * up`` -> ``fp8_linear`` (which
    re-quantises its input per row).  The SwiGLU rows are held in bf16 (the reference's activations are bf16 tensors)."""
    x_q, x_s = x_preq if x_preq is not None else quantize_rows_fp8(x)
    gu = scaled_mm(x_q, x_s, w13_q, w13_s)
    inter = w13_q.shape[0] // 2
    h = round_bf16(_silu(gu[:, :inter]) * gu[:, inter:])
    h_q, h_s = quantize_rows_fp8(h)
    return scaled_mm(h_q, h_s, w2_q, w2_s)


def quantize_moe_weights_fp8(w: MoEWeights) -> MoEWeightsInt8:
    """The materialised ``(out,in)`` forms of ``quantize_moe_weights`` quantised with ``quantize_fp8`` (per out-row);
    the ``*_q`` fields hold the E4M3 VALUES as fp32 (``e4m3_encode`` gives the bytes)."""
    e = w.gate.shape[0]
    w13 = np.concatenate([np.transpose(w.gate, (0, 2, 1)), np.transpose(w.up, (0, 2, 1))], axis=1)
    w2 = np.transpose(w.down, (0, 2, 1))
    q13, s13, q2, s2 = [], [], [], []
    for i in range(e):
        q, s = quantize_rows_fp8(w13[i]); q13.append(q); s13.append(s[:, 0])
        q, s = quantize_rows_fp8(w2[i]); q2.append(q); s2.append(s[:, 0])
    out = MoEWeightsInt8(np.stack(q13) if e else None, np.stack(s13) if e else None, np.stack(q2) if e else None,
                         np.stack(s2) if e else None)
    if w.shared_gate is not None:
        q, s_ = quantize_rows_fp8(np.concatenate([w.shared_gate, w.shared_up], axis=0)); out.sh13_q, out.sh13_s = q, s_[:, 0]
        q, s_ = quantize_rows_fp8(w.shared_down); out.sh2_q, out.sh2_s = q, s_[:, 0]
    return out


def moe_block_fp8(x: np.ndarray, token_ids: Optional[np.ndarray], table: np.ndarray, wq: MoEWeightsInt8,
                  cfg: MoEConfig) -> np.ndarray:
    """a3 with every expert (routed and shared) evaluated by :func:`fp8_swiglu`; the activations are quantised once per
    token (the same row serves both routes and the shared expert, as ``x_preq`` does for W8A8)."""
    n = x.shape[0]
    x_q, x_s = quantize_rows_fp8(x)
    routed = []
    if cfg.num_experts > 0:
        routes = route_table(token_ids, table, n, cfg.num_experts)
        for r in range(cfg.top_k):
            y = np.zeros((n, cfg.hidden_size), dtype=np.float32)
            for e in range(cfg.num_experts):
                rows = np.nonzero(routes[r] == e)[0]
                if rows.size:
                    y[rows] = fp8_swiglu(None, wq.w13_q[e], wq.w13_s[e], wq.w2_q[e], wq.w2_s[e], x_preq=(x_q[rows], x_s[rows]))
            routed.append(y)
    shared = fp8_swiglu(None, wq.sh13_q, wq.sh13_s, wq.sh2_q, wq.sh2_s, x_preq=(x_q, x_s)) if wq.sh13_q is not None else None
    if not routed:
        return shared
    return combine_fp32(routed, shared, cfg)


# --------------------------------------------------------------------------- #
# a11 / a13 — weight-only INT8 forms
# --------------------------------------------------------------------------- #

def quantize_experts_krow(gate_up: np.ndarray, down: np.ndarray):
    """``quantize_experts`` (core/quantization.py:424-454): per expert, ``quantize_int8``
    applied to ``gate_up[e] (H,2I)`` / ``down[e] (I,H)`` as stored, i.e. one scale per
    K-row (``gate_up_scale [E,H]``, ``down_scale [E,I]``)."""
    gq, gs, dq, ds = [], [], [], []
    for e in range(gate_up.shape[0]):
        q, s = quantize_rows_int8(gate_up[e]); gq.append(q); gs.append(s)
        q, s = quantize_rows_int8(down[e]); dq.append(q); ds.append(s)
    return np.stack(gq), np.stack(gs), np.stack(dq), np.stack(ds)


def fused_token_routed_int8_fp32(x, gu_q, gu_s, d_q, d_s, expert_ids, num_experts):
    """``fused_token_routed_forward_int8`` (kernels/fused_experts.py:153-291): weights
    dequantised (``int8 * scale[row]``) then the float path."""
    gu = gu_q.astype(np.float32) * gu_s[..., None].astype(np.float32)
    dn = d_q.astype(np.float32) * d_s[..., None].astype(np.float32)
    return fused_token_routed_fp32(x, gu, dn, expert_ids, num_experts)


def gemm_dequant_int8(x: np.ndarray, w_q: np.ndarray, w_s: np.ndarray) -> np.ndarray:
    """``I64_gemm_dequant_int8`` (csrc/I64_gemm.cu:40-78): ``Σ_k x[m,k]·(w[n,k]·s[n])`` fp32."""
    return _mm(x, (w_q.astype(np.float32) * np.asarray(w_s, np.float32)[:, None]).T)


def gemm_silu_int8(x, g_q, g_s, u_q, u_s) -> np.ndarray:
    """``I64_gemm_silu_int8`` (csrc/I64_gemm.cu:94-126)."""
    return _silu(gemm_dequant_int8(x, g_q, g_s)) * gemm_dequant_int8(x, u_q, u_s)


# --------------------------------------------------------------------------- #
# a14 — expert parallel, single-process "virtual ranks"
# --------------------------------------------------------------------------- #

def ep_forward_virtual(x_per_rank: Sequence[np.ndarray], ids_per_rank: Sequence[np.ndarray],
                       gate_up: np.ndarray, down: np.ndarray, num_experts: int) -> List[np.ndarray]:
    """``TokenRoutedMLP._forward_ep`` (layers/token_routed_i64/layer.py:209-298) for all
    ranks of an EP group at once.  dest rank = ``id // local_E`` (:227); tokens are
    stably sorted by destination rank (:241-243); received in source-rank order
    (all_to_all_single semantics, :253); local id = global − rank·local_E (:269);
    results return along the same splits (:286) and are un-sorted (:294-295)."""
    g = len(x_per_rank)
    local_e = num_experts // g
    order, sends = [], []
    for r in range(g):
        dest = np.asarray(ids_per_rank[r], np.int64) // local_e
        o = np.argsort(dest, kind="stable")
        order.append(o)
        sends.append([o[dest[o] == d] for d in range(g)])
    outs = [np.zeros_like(x_per_rank[r], dtype=np.float32) for r in range(g)]
    for d in range(g):
        recv_x = np.concatenate([x_per_rank[s][sends[s][d]] for s in range(g)], axis=0)
        recv_id = np.concatenate([np.asarray(ids_per_rank[s], np.int64)[sends[s][d]] for s in range(g)])
        local = recv_id - d * local_e
        y = fused_token_routed_fp32(recv_x, gate_up[d * local_e:(d + 1) * local_e],
                                    down[d * local_e:(d + 1) * local_e], local, local_e)
        pos = 0
        for s in range(g):
            cnt = len(sends[s][d])
            outs[s][sends[s][d]] = y[pos:pos + cnt]
            pos += cnt
    return outs


# --------------------------------------------------------------------------- #
# f1 — residual add + RMSNorm (+ activation quant): the step in front of the block
# --------------------------------------------------------------------------- #

def add_rmsnorm(x: np.ndarray, residual: Optional[np.ndarray], weight: np.ndarray, eps: float,
                bf16: bool) -> Tuple[np.ndarray, np.ndarray]:
    """``hidden = residual + hidden; hidden = post_attention_layernorm(hidden)`` (models/complexity_deep/model.py:358-361)
    with the eager ``RMSNorm.forward`` (layers/rmsnorm.py:86-90):
    ``norm = x.float().pow(2).mean(-1).add(eps).rsqrt(); (x.float()*norm).type_as(x) * weight``.
    ``bf16``: inputs hold bf16 values and every tensor the reference materialises in bf16 is rounded here (the sum, the
    ``type_as`` and the weight product); otherwise plain fp32 (also the contract of csrc/I64_rmsnorm.cu:98-190:
    ``x*rrms*weight``).  The row mean is taken in fp64 and rounded to fp32: the reference's fp32 reduction order is
    unspecified, so parity on the normalised rows is "equal up to one unit in the last place on isolated elements".
    Returns (new residual h, normalised rows y)."""
    x = np.asarray(x, np.float32)
    h = x if residual is None else (np.asarray(residual, np.float32) + x).astype(np.float32)
    if bf16 and residual is not None:
        h = round_bf16(h)
    ms = (h.astype(np.float64) ** 2).mean(axis=-1, keepdims=True).astype(np.float32)
    r = (1.0 / np.sqrt((ms + np.float32(eps)).astype(np.float64))).astype(np.float32)
    n1 = (h * r).astype(np.float32)
    w = np.asarray(weight, np.float32)
    if bf16:
        return h, round_bf16((round_bf16(n1) * w).astype(np.float32))
    return h, (n1 * w).astype(np.float32)


# --------------------------------------------------------------------------- #
# tolerance metric used by every floating-point parity test
# --------------------------------------------------------------------------- #

def rel_errors(got: np.ndarray, ref: np.ndarray) -> Tuple[float, float]:
    """(max-rel, mean-rel): max|got-ref| / max|ref| and mean|got-ref| / mean|ref|.
    north_star bar: ≤1e-2 and ≤1e-3."""
    got = np.asarray(got, np.float64)
    ref = np.asarray(ref, np.float64)
    d = np.abs(got - ref)
    return float(d.max() / max(np.abs(ref).max(), 1e-30)), float(d.mean() / max(np.abs(ref).mean(), 1e-30))
