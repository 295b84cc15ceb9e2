"""CPU oracle for the convex-MPC hot path — TEST INFRASTRUCTURE ONLY (parity unpinned).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this package.  See oracle/lmpc_oracle.h.
"""
