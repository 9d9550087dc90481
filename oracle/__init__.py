"""oracle/ -- TEST INFRASTRUCTURE ONLY (see oracle/ggml_ref.c header)."""
