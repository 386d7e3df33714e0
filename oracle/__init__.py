"""CPU oracle package — TEST INFRASTRUCTURE ONLY (see oracle/kgwe_oracle.h)."""
