"""Parallel state for the MoE FFN path: tensor-parallel state the layer reads, expert-parallel exchange."""
