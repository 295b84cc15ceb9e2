#!/bin/bash
export LMPC_DEBUG_STATS=1
run() { echo "== $*"; env "$@" python tools_profile.py 32768 northstar_a1 2 2>&1 | grep "lmpc\]" | tail -1; }
run LMPC_STEP_FRAC=0.995
run LMPC_STEP_FRAC=0.99
run LMPC_STEP_FRAC=0.999
run LMPC_STEP_FRAC=0.9999
run LMPC_SIGMA_POW=2
run LMPC_SIGMA_POW=4
run LMPC_SIGMA_POW=4 LMPC_STEP_FRAC=0.999
run LMPC_SIGMA_POW=2 LMPC_STEP_FRAC=0.999
