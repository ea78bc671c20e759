"""CPU oracle (test infrastructure only) — see oracle/dgs_oracle.c and oracle/field.py."""
