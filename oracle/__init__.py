"""CPU oracle (test infrastructure only; see lr_oracle.py's header)."""
