"""CPU oracle (test infrastructure only; see oracle/ctr_oracle.c)."""
