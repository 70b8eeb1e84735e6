"""TEST INFRASTRUCTURE ONLY — CPU checkers for the CUDA hot path (see oracle/pv3_oracle.c, oracle/ref_driver.cc).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this package."""
