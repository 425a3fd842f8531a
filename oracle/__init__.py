"""CPU oracle for the CCSD / (T) hot path: test infrastructure only (see ccsd_oracle.py header)."""
