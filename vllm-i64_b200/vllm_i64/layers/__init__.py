"""FFN layers of the hot path with the reference's constructor / method / state-dict names."""

from vllm_i64 import _overlay as _ov  # noqa: E402

_ov.extend(__name__, globals())
