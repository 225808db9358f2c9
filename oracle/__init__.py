"""CPU oracle (test infrastructure only - see oracle/cpu_odom.cpp header)."""
