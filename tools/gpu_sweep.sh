#!/bin/bash
# Solver-tuning sweep on a fixed 32768-problem workload (stats on stderr via LMPC_DEBUG_STATS).
export LMPC_DEBUG_STATS=1
run() { echo "== $*"; env "$@" python tools_profile.py 32768 northstar_a1 2 2>&1 | tail -2; }
run LMPC_TOL0=1e-7
run LMPC_TOL0=1e-7 LMPC_RDTOL0=1e-6
run LMPC_TOL0=1e-5 LMPC_RDTOL0=1e-6
run LMPC_TOL0=1e-4 LMPC_RDTOL0=1e-5
run LMPC_TOL0=1e-3 LMPC_RDTOL0=1e-4
run LMPC_TOL0=1e-7 LMPC_FZ0=0
run LMPC_TOL0=1e-7 LMPC_FZ0=0.3
run LMPC_TOL0=1e-5 LMPC_RDTOL0=1e-6 LMPC_FZ0=0
run LMPC_TOL0=1e-4 LMPC_RDTOL0=1e-5 LMPC_FZ0=0
