"""FFN layers of the hot path with the reference's constructor / method / state-dict names."""
