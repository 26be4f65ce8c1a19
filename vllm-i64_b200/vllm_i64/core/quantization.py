"""W8A8 quantisation ops — the surface of core/quantization.py:80-286 on the sm_100a kernels.

Semantics (bit-exact with the reference): per-row symmetric INT8 with ``s = max(absmax,1e-8)/127`` and
``q = clamp(round_half_even(x/s))`` in IEEE fp32; ``acc_i32 = q_x · q_wᵀ`` exact (tcgen05 ``kind::i8``);
``y = acc.float() * (s_x[:,None] * s_w[None,:])``.  The reference's four-tier fallback ladder
(``torch._int_mm`` → Triton → naive CUDA → dequant+F.linear, :157-224) is gone: one kernel, or an exception.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional, Tuple

import torch

from vllm_i64 import _C


@dataclass
class QuantConfig:
    method: str = "none"   # "none" | "int8"
    group_size: int = 128


def _quantize_rows(x2d: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    _C.require_cuda(x2d)
    if x2d.dtype not in (torch.float32, torch.bfloat16, torch.float16):
        x2d = x2d.float()
    x2d = x2d.contiguous()
    rows, cols = x2d.shape
    q = torch.empty(rows, cols, dtype=torch.int8, device=x2d.device)
    s = torch.empty(rows, dtype=torch.float32, device=x2d.device)
    _C.check(_C.lib().i64b_quantize_rows_int8(x2d.data_ptr(), _C.dtype_code(x2d), q.data_ptr(), s.data_ptr(), rows, cols,
                                              _C.stream_ptr(x2d.device)))
    return q, s


def quantize_int8(weight: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """Per-channel (per out-row) symmetric INT8: ``(out,in)`` → int8 + ``(out,)`` fp32 scale
    (core/quantization.py:80-91).  Arithmetic is fp32 whatever the storage dtype.  Weights still on the HOST (the reference's
    loader quantises before it moves the model, core/loader.py:627-651) go through the reference's own torch ops when this
    package overlays it — load time only, never the forward path."""
    if not weight.is_cuda and _REF_MOD is not None:
        return _REF_MOD.quantize_int8(weight)
    shape = weight.shape
    q, s = _quantize_rows(weight.reshape(-1, shape[-1]))
    return q.view(shape), s.view(shape[:-1])


def quantize_activations_int8(x: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """Dynamic per-token INT8 (core/quantization.py:103-116)."""
    shape = x.shape
    q, s = _quantize_rows(x.reshape(-1, shape[-1]))
    return q.view(shape), s.view(shape[:-1])


def dequantize_int8(quantized: torch.Tensor, scale: torch.Tensor) -> torch.Tensor:
    """``int8 * scale[row]`` → fp32 (core/quantization.py:94-96)."""
    _C.require_cuda(quantized, scale)
    shape = quantized.shape
    q = quantized.reshape(-1, shape[-1]).contiguous()
    s = scale.reshape(-1).float().contiguous()
    out = torch.empty(q.shape, dtype=torch.float32, device=q.device)
    _C.check(_C.lib().i64b_dequantize_int8(q.data_ptr(), s.data_ptr(), out.data_ptr(), q.shape[0], q.shape[1],
                                           _C.stream_ptr(q.device)))
    return out.view(shape)


def int8_gemm(x_int8: torch.Tensor, x_scale: Optional[torch.Tensor], w_int8: torch.Tensor,
              w_scale: Optional[torch.Tensor], want_acc: bool = False):
    """``(M,K) int8 · (N,K) int8ᵀ`` → fp32 ``(M,N)`` rescaled, and optionally the raw INT32 accumulators."""
    _C.require_cuda(x_int8, w_int8)
    a = x_int8.contiguous()
    b = w_int8.contiguous()
    m, k = a.shape
    n = b.shape[0]
    out = None
    if x_scale is not None and w_scale is not None:
        out = torch.empty(m, n, dtype=torch.float32, device=a.device)
        xs, wsc = x_scale.float().contiguous(), w_scale.float().contiguous()
    acc = torch.empty(m, n, dtype=torch.int32, device=a.device) if want_acc else None
    _C.check(_C.lib().i64b_int8_gemm(a.data_ptr(), _C.ptr(xs) if out is not None else None, b.data_ptr(),
                                     _C.ptr(wsc) if out is not None else None, _C.ptr(out), _C.ptr(acc), m, n, k,
                                     _C.stream_ptr(a.device)))
    return (out, acc) if want_acc else out


def int8_linear_native(x: torch.Tensor, weight_int8: torch.Tensor, weight_scale: torch.Tensor,
                       bias: Optional[torch.Tensor] = None,
                       x_preq: Optional[Tuple[torch.Tensor, torch.Tensor]] = None) -> torch.Tensor:
    """``y = x @ Wᵀ`` in W8A8 (core/quantization.py:119-176).  Returns fp32 like the reference."""
    orig = x.shape
    x2 = x.reshape(-1, orig[-1])
    if x_preq is not None:
        xq, xs = x_preq[0].reshape(-1, orig[-1]), x_preq[1].reshape(-1)
    else:
        xq, xs = quantize_activations_int8(x2)
    out = int8_gemm(xq, xs, weight_int8, weight_scale)
    if bias is not None:
        out = out + bias
    return out.reshape(*orig[:-1], weight_int8.shape[0])


def int8_fused_gate_up_native(x: torch.Tensor, fused_int8: torch.Tensor, fused_scale: torch.Tensor, inter_size: int,
                              x_preq: Optional[Tuple[torch.Tensor, torch.Tensor]] = None
                              ) -> Tuple[torch.Tensor, torch.Tensor]:
    """One activation quantisation + one INT8 GEMM for gate|up (core/quantization.py:227-286)."""
    res = int8_linear_native(x, fused_int8, fused_scale, x_preq=x_preq)
    gate, up = res.split(inter_size, dim=-1)
    return gate, up


def int8_linear_available(device: str = "cuda") -> bool:
    return device != "cpu" and torch.cuda.is_available()


def quantize_experts(gate_up: torch.Tensor, down: torch.Tensor, config: QuantConfig) -> dict:
    """Per-expert weight-only quantisation with one scale per K-row (core/quantization.py:424-454)."""
    if config.method == "none":
        return {"gate_up": gate_up, "down": down, "method": "none"}
    if config.method != "int8":
        raise ValueError(f"unsupported quantisation method {config.method!r} on this path (int8 only)")
    gq, gs = quantize_int8(gate_up)
    dq, ds = quantize_int8(down)
    return {"method": "int8", "gate_up_int8": gq, "gate_up_scale": gs, "down_int8": dq, "down_scale": ds}


# names the reference's module has and this one does not (INT4 helpers, capability flags): taken from the reference when
# this package runs as an overlay on it; load-time helpers called on CPU tensors delegate to the reference's own code
from vllm_i64 import _overlay as _ov  # noqa: E402

_REF_MOD = _ov.fill(__name__, globals())
