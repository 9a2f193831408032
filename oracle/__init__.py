"""CPU oracle (test infrastructure only -- see dgs_oracle.py header)."""
