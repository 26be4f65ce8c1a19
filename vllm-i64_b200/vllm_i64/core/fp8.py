"""FP8 E4M3 ops — the surface of core/fp8.py:54-290 on the sm_100a kernels.

Semantics of the reference's GPU branch: weights quantised per out-row and activations per token row with
``scale = max(absmax, 1e-12) / 448`` and ``q = e4m3((x / scale).clamp(-448, 448))`` in IEEE fp32 (bit-exact here, see
tests/golden/tiny_fp8.npz), every projection ``y = (x_q . w_q^T) * (x_scale[:, None] * w_scale[None, :])`` with exact E4M3
products and fp32 accumulation — the reference's ``torch._scaled_mm`` call becomes one tcgen05 ``kind::f8f6f4`` GEMM
(``i64b_fp8_gemm``).  The reference's hardware ladder (``fp8_available()`` → INT8 on SM80 → dequantise + ``F.linear``
elsewhere) is gone: CUDA tensors run the kernel; CPU tensors raise, except under the overlay (the reference package importable beside this one), where
the load-time and host-tensor calls are handed to the reference's own module.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from vllm_i64 import _C

_FP8_DTYPE = torch.float8_e4m3fn
_FP8_MAX = 448.0


def fp8_available() -> bool:
    """FP8 compute needs the sm_100a library and a GPU (core/fp8.py:54-65 asks for SM89+ and ``torch._scaled_mm``)."""
    return torch.cuda.is_available()


def fp8_dtype():
    return _FP8_DTYPE


def _quantize_rows(x2d: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    _C.require_cuda(x2d)
    if x2d.dtype not in (torch.float32, torch.bfloat16, torch.float16):
        x2d = x2d.float()
    x2d = x2d.contiguous()
    rows, cols = x2d.shape
    if cols % 2:
        raise ValueError("FP8 row quantisation needs an even number of columns")
    q = torch.empty(rows, cols, dtype=torch.uint8, device=x2d.device)
    s = torch.empty(rows, dtype=torch.float32, device=x2d.device)
    _C.check(_C.lib().i64b_quantize_rows_fp8(x2d.data_ptr(), _C.dtype_code(x2d), q.data_ptr(), s.data_ptr(), rows, cols,
                                             _C.stream_ptr(x2d.device)))
    return q.view(_FP8_DTYPE), s


def quantize_fp8(weight: torch.Tensor, per_channel: bool = True) -> Tuple[torch.Tensor, torch.Tensor]:
    """``(out, in)`` float → ``float8_e4m3fn`` + ``(out,)`` fp32 scale, or one scalar scale (core/fp8.py:75-110)."""
    if not weight.is_cuda and _REF_MOD is not None:
        return _REF_MOD.quantize_fp8(weight, per_channel)      # load time, weights still on the host: the reference's own torch ops
    if per_channel:
        shape = weight.shape
        q, s = _quantize_rows(weight.reshape(-1, shape[-1]))
        return q.view(shape), s.view(shape[:-1])
    _C.require_cuda(weight)
    q, s = _quantize_rows(weight.reshape(1, -1))          # one row = the whole tensor: the per-tensor rule
    return q.view(weight.shape), s.reshape(())


def quantize_activations_fp8(x: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """Dynamic per-token FP8: ``(tokens, features)`` → fp8 + ``(tokens, 1)`` scale (core/fp8.py:113-129)."""
    shape = x.shape
    q, s = _quantize_rows(x.reshape(-1, shape[-1]))
    return q.view(shape), s.view(*shape[:-1], 1)


def fp8_gemm(x_fp8: torch.Tensor, x_scale: torch.Tensor, w_fp8: torch.Tensor, w_scale: torch.Tensor,
             out_dtype: torch.dtype = torch.float32) -> torch.Tensor:
    """``(M,K) e4m3 · (N,K) e4m3ᵀ`` with per-row × per-column scales → ``(M,N)`` fp32 or bf16."""
    _C.require_cuda(x_fp8, w_fp8, x_scale, w_scale)
    a = x_fp8.contiguous().view(torch.uint8)
    b = w_fp8.contiguous().view(torch.uint8)
    m, k = a.shape
    n = b.shape[0]
    odt = torch.bfloat16 if out_dtype == torch.bfloat16 else torch.float32
    out = torch.empty(m, n, dtype=odt, device=a.device)
    xs, ws = x_scale.reshape(-1).float().contiguous(), w_scale.reshape(-1).float().contiguous()
    if xs.numel() != m or ws.numel() != n:
        raise RuntimeError("fp8_gemm needs one scale per row of x and one per row of w")
    _C.check(_C.lib().i64b_fp8_gemm(a.data_ptr(), xs.data_ptr(), b.data_ptr(), ws.data_ptr(), out.data_ptr(),
                                    _C.dtype_code(out), m, n, k, _C.stream_ptr(a.device)))
    return out if out.dtype == out_dtype else out.to(out_dtype)


def fp8_linear(x: torch.Tensor, weight_fp8: torch.Tensor, weight_scale: torch.Tensor,
               bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``y = x @ Wᵀ`` in FP8 (core/fp8.py:137-193): quantise the activations per token, one scaled GEMM, output in the
    dtype of ``x`` (fp16 for non-float inputs, as the reference)."""
    if not x.is_cuda and _REF_MOD is not None:
        return _REF_MOD.fp8_linear(x, weight_fp8, weight_scale, bias)      # host tensors: the reference's dequantise + F.linear
    orig = x.shape
    x2 = x.reshape(-1, orig[-1])
    xq, xs = quantize_activations_fp8(x2)
    out = fp8_gemm(xq, xs, weight_fp8, weight_scale, x.dtype if x.is_floating_point() else torch.float16)
    if bias is not None:
        out = out + bias
    return out.reshape(*orig[:-1], weight_fp8.shape[0])


def fp8_fused_gate_up(x: torch.Tensor, fused_fp8: torch.Tensor, fused_scale: torch.Tensor, inter_size: int
                      ) -> Tuple[torch.Tensor, torch.Tensor]:
    """One activation quantisation + one FP8 GEMM for ``cat([gate, up])`` (core/fp8.py:196-245)."""
    if not x.is_cuda and _REF_MOD is not None:
        return _REF_MOD.fp8_fused_gate_up(x, fused_fp8, fused_scale, inter_size)
    res = fp8_linear(x, fused_fp8, fused_scale)
    gate, up = res.split(inter_size, dim=-1)
    return gate, up


def quantize_experts_fp8(gate_up: torch.Tensor, down: torch.Tensor) -> dict:
    """All experts at once, one scale per row of the stored ``[E, H, 2I]`` / ``[E, I, H]`` tensors (core/fp8.py:252-277)."""
    gq, gs = quantize_fp8(gate_up)
    dq, ds = quantize_fp8(down)
    return {"method": "fp8", "gate_up_fp8": gq, "gate_up_scale": gs, "down_fp8": dq, "down_scale": ds}


def quantize_dense_fp8(weight: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """One dense ``(out, in)`` matrix (core/fp8.py:284-290)."""
    return quantize_fp8(weight, per_channel=True)


# anything else the reference's module has is taken from it when this package runs as an overlay
from vllm_i64 import _overlay as _ov  # noqa: E402

_REF_MOD = _ov.fill(__name__, globals())
