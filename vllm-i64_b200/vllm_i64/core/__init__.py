"""Hot-path slice of ``vllm_i64.core``: the W8A8 quantisation ops."""
