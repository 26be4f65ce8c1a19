"""Synthetic stand-ins for checkpoint data that the reference does not ship (numpy only; shared by the parity
tests, the oracle and bench.py so that every arm sees the same tables)."""
import numpy as np


def synthetic_hash_table(vocab_size: int, num_experts: int, top_k: int, layer: int) -> np.ndarray:
    """Layer-specific balanced hash table ``[K, V]`` standing in for the checkpoint's ``topk_token_to_expert``
    (SURVEY §8(d)); the real tables ship in the checkpoint (core/loader.py:374-387).  Route 0 is an exactly
    balanced pseudo-random assignment; route r ≥ 1 is offset by a token- and layer-dependent step in [1, E-1], so
    it never collides with route 0."""
    rng = np.random.RandomState(1000 + layer)
    perm = rng.permutation(vocab_size)
    t0 = np.empty(vocab_size, dtype=np.int64)
    t0[perm] = np.arange(vocab_size, dtype=np.int64) % num_experts
    rows = [t0]
    v = np.arange(vocab_size, dtype=np.int64)
    for r in range(1, top_k):
        h = (v * 2654435761 + (layer + 1) * 40503 + r * 977) % (2 ** 32)
        step = 1 + (h >> 7) % max(1, num_experts - 1)
        rows.append((t0 + step) % num_experts if num_experts > 1 else t0.copy())
    return np.stack(rows)
