"""CPU oracle (test infrastructure only; see rnnt_oracle.py / rnnt_oracle.c headers)."""
