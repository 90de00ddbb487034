"""CPU oracle package -- test infrastructure only (see bert_oracle.c)."""
