"""vllm_i64 — B200-native (sm_100a) drop-in for the token-routed MoE FFN path of Complexity-ML/vllm-i64.

Only the hot path lives here: ``kernels`` (binding shim over the C ABI in ``include/i64b.h`` — no Triton, no
multi-backend dispatch, no CPU fallback), ``layers`` (``TokenRoutedMLP`` / ``DenseMLP`` with the reference's
constructor, method and state-dict names), ``core.quantization`` (the W8A8 ops) and ``parallel`` (TP state the
layer reads + the expert-parallel exchange).  Everything else of the reference (engine, scheduler, KV cache,
server) runs unchanged on top of these modules.
"""
__version__ = "0.1.0"

# overlay on an installed reference package of the same name, when there is one (see _overlay.py)
from vllm_i64 import _overlay as _ov  # noqa: E402

_ov.extend(__name__, globals())
patch_reference_models = _ov.patch_reference_models
collect_expert_load = _ov.collect_expert_load
