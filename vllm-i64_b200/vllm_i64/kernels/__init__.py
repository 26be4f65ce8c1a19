"""Binding shim over the sm_100a C ABI (include/i64b.h): same names and signatures as the reference's
``vllm_i64.kernels`` package; no Triton, no backend ladder, no CPU fallback."""
from vllm_i64.kernels.kernel_loader import get_ops, is_cuda_kernels_available  # noqa: F401

from vllm_i64 import _overlay as _ov  # noqa: E402

_ov.extend(__name__, globals())
