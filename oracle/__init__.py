"""CPU oracle for the APM hot path - TEST INFRASTRUCTURE ONLY (see oracle/apm_oracle.cpp header)."""
