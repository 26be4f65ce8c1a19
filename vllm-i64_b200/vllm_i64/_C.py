"""ctypes binding of ``libi64b.so`` (the C ABI declared in ``include/i64b.h``).

The product path has no fallback: if the shared library is missing or a call fails, an exception is raised.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(os.path.dirname(_HERE), "lib", "libi64b.so")

BF16, F32, F16 = 0, 1, 2
STAGE_PLAN, STAGE_DISPATCH, STAGE_GEMM1, STAGE_REQUANT, STAGE_GEMM2, STAGE_COMBINE, STAGE_ALL = 1, 2, 4, 8, 16, 32, 0xFF
MODE_BF16, MODE_W8A8, MODE_FP8 = 0, 1, 2
DEBUG_KEEP_TAPS, DEBUG_FUSE_COMBINE, DEBUG_NO_FUSED, DEBUG_TRI, DEBUG_NO_TRI, DEBUG_TRI_FOLD = 1, 2, 4, 8, 16, 32
DEBUG_NO_OVERLAP, DEBUG_GATHER = 64, 128
TRI_MIN_ROWS_ONEACC, TRI_MIN_ROWS_FOLD = 288, 512      # include/i64b.h
(TAP_ROUTES, TAP_DEST_ROW, TAP_SCATTER_IDX, TAP_COUNTS, TAP_OFFSETS, TAP_XS, TAP_XS_SCALE, TAP_INTER_R, TAP_INTER_S,
 TAP_INTER_SCALE_R, TAP_INTER_SCALE_S, TAP_Q14_R, TAP_Q14_S, TAP_Y, TAP_GROUP_START) = range(15)

_DTYPES = {torch.bfloat16: BF16, torch.float32: F32, torch.float16: F16}


class I64bError(RuntimeError):
    pass


class MoEConfig(C.Structure):
    _fields_ = [("hidden_size", C.c_int32), ("num_experts", C.c_int32), ("expert_inter", C.c_int32),
                ("shared_inter", C.c_int32), ("vocab_size", C.c_int32), ("top_k", C.c_int32), ("mode", C.c_int32),
                ("reserved", C.c_int32), ("route_weight", C.c_float * 4), ("routed_scale", C.c_float),
                ("shared_scale", C.c_float)]


EP_MAX_RANKS = 8
EP_FLAGS_INTS = 1024


class EPCtx(C.Structure):
    """``i64b_ep_ctx`` (include/i64b.h): peer-mapped buffers of every rank of the expert-parallel group."""
    _fields_ = [("rank", C.c_int32), ("size", C.c_int32), ("n_local", C.c_int32), ("hidden", C.c_int32),
                ("top_k", C.c_int32), ("num_experts", C.c_int32), ("vocab", C.c_int32), ("y_dtype", C.c_int32),
                ("x_mode", C.c_int32), ("comm_ctas", C.c_int32),
                ("cap_src", C.c_int64), ("peer_xs", C.c_void_p * EP_MAX_RANKS),
                ("peer_xs_scale", C.c_void_p * EP_MAX_RANKS), ("peer_y", C.c_void_p * EP_MAX_RANKS),
                ("peer_flags", C.c_void_p * EP_MAX_RANKS), ("state", C.c_void_p)]


_lib: Optional[C.CDLL] = None

_P, _I, _L, _Z, _F = C.c_void_p, C.c_int, C.c_int64, C.c_size_t, C.c_float
_SIGNATURES = {
    "i64b_last_error": (C.c_char_p, []),
    "i64b_version": (_I, []),
    "i64b_launch_count": (_L, []),
    "i64b_launch_count_reset": (None, []),
    "i64b_silu_lut_q7": (C.POINTER(C.c_int16), []),
    "i64b_route_modulo": (_I, [_P, _P, _L, _I, _P]),
    "i64b_pack_table": (_I, [_P, _P, _I, _I, _I, _P]),
    "i64b_route_table": (_I, [_P, _P, _P, _L, _I, _I, _I, _P]),
    "i64b_plan_workspace_bytes": (_Z, [_L, _I, _I]),
    "i64b_dispatch_plan": (_I, [_P, _I, _L, _I, _P, _P, _P, _P, _P, _P, _Z, _P]),
    "i64b_dispatch_plan_static": (_I, [_P, _I, _L, _I, _P, _P, _P, _P, _P, _P]),     # scratch inside the library
    "i64b_scatter_rows": (_I, [_P, _P, _P, _L, _L, _P]),
    "i64b_gather_rows": (_I, [_P, _P, _P, _L, _L, _P]),
    "i64b_combine": (_I, [_P, _P, _P, _I, C.POINTER(_F), _F, _F, _P, _I, _L, _I, _P]),
    "i64b_silu_hadamard": (_I, [_P, _P, _L, _I, _I, _P]),
    "i64b_quantize_rows_int8": (_I, [_P, _I, _P, _P, _L, _L, _P]),
    "i64b_dequantize_int8": (_I, [_P, _P, _P, _L, _L, _P]),
    "i64b_add_rmsnorm": (_I, [_P, _P, _P, _I, _F, _P, _P, _P, _P, _L, _I, _P]),
    "i64b_moe_create": (_I, [C.POINTER(MoEConfig), C.POINTER(_P)]),
    "i64b_moe_destroy": (None, [_P]),
    "i64b_moe_set_scales": (_I, [_P, C.POINTER(_F), _I, _F, _F]),
    "i64b_moe_packed_bytes": (_Z, [_P]),
    "i64b_moe_load_bf16": (_I, [_P, _P, _P, _P, _P, _P, _P, _P, _P]),
    "i64b_moe_load_w8a8": (_I, [_P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P]),
    "i64b_quantize_rows_fp8": (_I, [_P, _I, _P, _P, _L, _L, _P]),
    "i64b_fp8_gemm": (_I, [_P, _P, _P, _P, _P, _I, _L, _L, _L, _P]),
    "i64b_moe_load_fp8": (_I, [_P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P]),
    "i64b_moe_set_table": (_I, [_P, _P]),
    "i64b_moe_workspace_bytes": (_Z, [_P, _L]),
    "i64b_moe_forward": (_I, [_P, _P, _P, _P, _I, _L, _P, _Z, _P]),
    "i64b_moe_forward_preq": (_I, [_P, _P, _P, _P, _P, _I, _L, _P, _Z, _P]),
    "i64b_moe_uses_fused": (_I, [_P, _L]),
    "i64b_moe_fuses_combine": (_I, [_P, _L]),
    "i64b_moe_tri_form": (_I, [_P, _L]),
    "i64b_moe_forward_stages": (_I, [_P, _P, _P, _P, _I, _L, _P, _Z, C.c_uint, _P]),
    "i64b_moe_forward_routed": (_I, [_P, _P, _P, _I, C.POINTER(_F), _F, _F, _I, _P, _I, _L, _P, _Z, _P]),
    "i64b_moe_workspace_bytes_routed": (_Z, [_P, _L, _I]),
    "i64b_copy_device": (_I, [_P, _P, _Z, _P]),
    "i64b_moe_forward_host": (_I, [_P, _P, _P, _P, _I, _L, _P, _P, _P, _P, _Z, _P]),
    "i64b_moe_tap": (_P, [_P, _I]),
    "i64b_moe_set_debug": (_I, [_P, _I]),
    "i64b_moe_set_next": (_I, [_P, _P]),
    "i64b_moe_set_cta_limit": (_I, [_P, _I]),
    "i64b_debug_timeline_dump": (_I, []),
    "i64b_fused_decode_supported": (_I, [_I, _I, _I, _I, _I, _I]),
    "i64b_ep_alloc": (_I, [_Z, C.POINTER(_P)]),
    "i64b_ep_free": (_I, [_P]),
    "i64b_ipc_export": (_I, [_P, C.c_char_p]),
    "i64b_ipc_import": (_I, [C.c_char_p, C.POINTER(_P)]),
    "i64b_ipc_close": (_I, [_P]),
    "i64b_ep_workspace_bytes": (_Z, [C.POINTER(EPCtx)]),
    "i64b_ep_dispatch": (_I, [C.POINTER(EPCtx), _P, _P, _P, _P, _Z, _P]),
    "i64b_ep_plan": (_I, [C.POINTER(EPCtx), _P, _P, _P, _Z, _P]),
    "i64b_ep_barrier": (_I, [C.POINTER(EPCtx), _P]),
    "i64b_ep_combine": (_I, [C.POINTER(EPCtx), _P, _P, C.POINTER(_F), _F, _F, _P, _I, _P]),
    "i64b_ep_signal": (_I, [C.POINTER(EPCtx), _P]),
    "i64b_ep_wait": (_I, [C.POINTER(EPCtx), _P]),
    "i64b_moe_dispatched_workspace_bytes": (_Z, [_P, C.POINTER(EPCtx)]),
    "i64b_moe_forward_dispatched": (_I, [_P, C.POINTER(EPCtx), _P, _P, _Z, _P]),
    "i64b_int8_gemm": (_I, [_P, _P, _P, _P, _P, _P, _L, _L, _L, _P]),
    "i64b_bf16_gemm": (_I, [_P, _P, _P, _P, _I, _L, _L, _L, _P]),
}
EXPORTED = tuple(sorted(_SIGNATURES))


def lib() -> C.CDLL:
    """Load (once) and return the C-ABI library.  Raises if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} not found: build the CUDA library first (python -c 'import __graft_entry__ as g; "
                "g.build()' or make -C vllm-i64_b200/csrc).  There is no CPU or PyTorch fallback for this path.")
        handle = C.CDLL(LIB_PATH)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(handle, name)
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


def check(rc: int) -> None:
    if rc != 0:
        raise I64bError(f"i64b error {rc}: {lib().i64b_last_error().decode(errors='replace')}")


def stream_ptr(device=None) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def dtype_code(t: torch.Tensor) -> int:
    try:
        return _DTYPES[t.dtype]
    except KeyError:
        raise TypeError(f"unsupported dtype {t.dtype}") from None


def require_cuda(*tensors: torch.Tensor) -> None:
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise RuntimeError(
                "vllm_i64 (B200 build): this op runs only on CUDA tensors; there is no CPU fallback on this path")


def launch_count() -> int:
    return int(lib().i64b_launch_count())


def launch_count_reset() -> None:
    lib().i64b_launch_count_reset()


_workspaces = {}


def workspace(device: torch.device, nbytes: int) -> torch.Tensor:
    """Per-(device, stream) scratch buffer, grown on demand; owned by torch's caching allocator."""
    key = (device.index, torch.cuda.current_stream(device).cuda_stream)
    buf = _workspaces.get(key)
    if buf is None or buf.numel() < nbytes:
        buf = torch.empty(int(nbytes * 1.25) + 4096, dtype=torch.uint8, device=device)
        _workspaces[key] = buf
    return buf


def tap_tensor(layer_handle, what: int, shape, dtype: torch.dtype, device) -> torch.Tensor:
    """View a debug tap of the last forward as a tensor (copy)."""
    p = lib().i64b_moe_tap(layer_handle, what)
    if not p:
        raise I64bError(f"tap {what} unavailable")
    n = 1
    for s in shape:
        n *= int(s)
    out = torch.empty(shape, dtype=dtype, device=device)
    if n:
        check(lib().i64b_copy_device(out.data_ptr(), p, n * out.element_size(), stream_ptr(out.device)))
    return out
