"""CPU oracle (test infrastructure only): see oracle/pf_oracle.h. Never imported by bachelors_b200."""
