"""Hot-path slice of ``vllm_i64.core``: the W8A8 quantisation ops."""

from vllm_i64 import _overlay as _ov  # noqa: E402

_ov.extend(__name__, globals())
