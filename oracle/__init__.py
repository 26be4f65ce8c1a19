"""CPU oracle (test infrastructure only — see oracle/i64_oracle.py)."""
