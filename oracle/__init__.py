"""CPU oracle (test infrastructure only; see oracle/flat.py and oracle/flat_oracle.c)."""
