"""``get_i64_cuda_ops()`` — the ``I64_*`` op module of csrc/I64_binding.cu:49-87 (FFN-side subset).

The reference JIT-compiles rmsnorm/rope/quantize/softmax/gemm into one extension
(kernels/cuda/I64_loader.py:58-84).  The FFN path needs quantize_int8, dequantize_int8,
quantize_perchannel_int8, gemm_dequant_int8 and gemm_silu_int8, and — the step in front of the block —
rmsnorm_forward / rmsnorm_quant_forward plus their fused residual-add form ``add_rmsnorm``; rope / softmax (attention
side) are out of scope.
"""
from __future__ import annotations

import torch

from vllm_i64 import _C
from vllm_i64.core import quantization as _Q


class _I64CudaOps:
    @staticmethod
    def quantize_int8(x: torch.Tensor):
        """Per-token INT8 of a 2-D tensor → (int8, fp32 scale) (csrc/I64_quantize.cu:173-197)."""
        if x.dim() != 2:
            raise RuntimeError("x must be 2D")
        return _Q.quantize_activations_int8(x)

    @staticmethod
    def dequantize_int8(x_int8: torch.Tensor, scale: torch.Tensor) -> torch.Tensor:
        return _Q.dequantize_int8(x_int8, scale)

    @staticmethod
    def quantize_perchannel_int8(weight: torch.Tensor):
        """Per out-row weight quantisation (csrc/I64_quantize.cu:223-243)."""
        if weight.dim() != 2:
            raise RuntimeError("weight must be 2D")
        return _Q.quantize_int8(weight)

    @staticmethod
    def add_rmsnorm(x: torch.Tensor, residual, weight: torch.Tensor, eps: float = 1e-6, quantize: bool = False,
                    want_normed: bool = True):
        """ONE pass: ``h = residual + x`` (skipped when ``residual`` is None), ``y = rmsnorm(h) * weight`` and, with
        ``quantize``, ``(q, scale) = quantize_activations_int8(y)`` (include/i64b.h, i64b_add_rmsnorm).  bf16 or fp32
        2-D rows.  Returns ``(h | None, y | None, (q, scale) | None)``."""
        _C.require_cuda(x, residual, weight)
        if x.dim() != 2:
            raise RuntimeError("x must be 2D")
        if x.dtype not in (torch.bfloat16, torch.float32):
            raise TypeError("add_rmsnorm: bf16 or fp32 rows")
        x = x.contiguous()
        w = weight.to(x.dtype).contiguous()
        res = None if residual is None else residual.to(x.dtype).contiguous()
        n, h = x.shape
        h_out = torch.empty_like(x) if res is not None else None
        y = torch.empty_like(x) if (want_normed or not quantize) else None
        q = torch.empty(n, h, dtype=torch.int8, device=x.device) if quantize else None
        sc = torch.empty(n, dtype=torch.float32, device=x.device) if quantize else None
        with torch.cuda.device(x.device):
            _C.check(_C.lib().i64b_add_rmsnorm(x.data_ptr(), _C.ptr(res), w.data_ptr(), _C.dtype_code(x), float(eps),
                                               _C.ptr(h_out), _C.ptr(y), _C.ptr(q), _C.ptr(sc), n, h, _C.stream_ptr(x.device)))
        return h_out, y, ((q, sc.unsqueeze(-1)) if quantize else None)

    @staticmethod
    def rmsnorm_forward(x: torch.Tensor, weight: torch.Tensor, eps: float = 1e-6) -> torch.Tensor:
        """``I64_rmsnorm_forward`` (csrc/I64_rmsnorm.cu:41-90, csrc/I64_binding.cu:49-87)."""
        return _I64CudaOps.add_rmsnorm(x, None, weight, eps)[1]

    @staticmethod
    def rmsnorm_quant_forward(x: torch.Tensor, weight: torch.Tensor, eps: float = 1e-6):
        """``I64_rmsnorm_quant_forward`` (csrc/I64_rmsnorm.cu:98-190): ``(int8 [N,H], scale [N,1])`` — the ``x_preq`` of
        layers/rmsnorm.py:29-54 / layers/dense_mlp.py:46-58."""
        return _I64CudaOps.add_rmsnorm(x, None, weight, eps, quantize=True, want_normed=False)[2]

    @staticmethod
    def gemm_dequant_int8(x: torch.Tensor, w_int8: torch.Tensor, w_scale: torch.Tensor) -> torch.Tensor:
        """``Σ_k x[m,k]·(w[n,k]·s[n])`` → fp32 (csrc/I64_gemm.cu:40-78).  The int8 weights are exact in bf16, the
        fp32 activations are split into bf16 hi+lo halves stacked along K, so one tcgen05 bf16 GEMM with the
        per-channel scale in its epilogue reproduces the fp32-activation product to ~2^-17 relative."""
        _C.require_cuda(x, w_int8, w_scale)
        x32 = x.float().contiguous()
        hi = x32.to(torch.bfloat16)
        lo = (x32 - hi.float()).to(torch.bfloat16)
        a = torch.cat([hi, lo], dim=1).contiguous()
        wb = w_int8.to(torch.bfloat16)
        b = torch.cat([wb, wb], dim=1).contiguous()
        m, k2 = a.shape
        n = b.shape[0]
        out = torch.empty(m, n, dtype=torch.float32, device=x.device)
        s = w_scale.float().contiguous()
        _C.check(_C.lib().i64b_bf16_gemm(a.data_ptr(), b.data_ptr(), s.data_ptr(), out.data_ptr(), _C.F32, m, n, k2,
                                         _C.stream_ptr(x.device)))
        return out

    @staticmethod
    def gemm_silu_int8(x, gate_int8, gate_scale, up_int8, up_scale) -> torch.Tensor:
        """``silu(x@Wgᵀ)·(x@Wuᵀ)`` with int8 weights (csrc/I64_gemm.cu:94-126)."""
        from vllm_i64.kernels.kernel_loader import get_ops
        w = torch.cat([gate_int8, up_int8], dim=0)
        s = torch.cat([gate_scale, up_scale], dim=0)
        gu = _I64CudaOps.gemm_dequant_int8(x, w, s)
        return get_ops().silu_hadamard(gu, gate_int8.shape[0])


_OPS = None


def get_i64_cuda_ops():
    """Returns the op object; raises ImportError if libi64b.so is missing (no silent ``None``)."""
    global _OPS
    if _OPS is None:
        _C.lib()
        _OPS = _I64CudaOps()
    return _OPS


def is_available() -> bool:
    try:
        _C.lib()
    except (ImportError, OSError):
        return False
    return torch.cuda.is_available()


# the reference's I64_loader module lives beside this file in its own tree (kernels/cuda/I64_loader.py)
from vllm_i64 import _overlay as _ov  # noqa: E402

if _ov.active():
    import os as _os
    _ref_cuda = _os.path.join(_ov._REF, "kernels", "cuda")
    if _os.path.isdir(_ref_cuda) and _ref_cuda not in __path__:
        __path__.append(_ref_cuda)
    is_i64_cuda_available = is_available
