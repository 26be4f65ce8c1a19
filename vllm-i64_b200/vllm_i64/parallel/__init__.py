"""Parallel state for the MoE FFN path: tensor-parallel state the layer reads, expert-parallel exchange."""

from vllm_i64 import _overlay as _ov  # noqa: E402

_ov.extend(__name__, globals())
