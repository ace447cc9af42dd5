"""CPU oracle for the CUDA engine — TEST INFRASTRUCTURE ONLY (see oracle/mg_oracle.c header)."""
