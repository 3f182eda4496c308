"""CPU oracle package -- TEST INFRASTRUCTURE ONLY (see oracle/oracle.c header).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this.  The product package mcmcpp_b200 never does."""
