#!/bin/bash
set -x
mkdir -p gpurun_out
make -C legged_mpc_control_b200/csrc > /dev/null 2>&1
python tools/gpu_sweep_fail.py > gpurun_out/sweep_fail.log 2>&1; cat gpurun_out/sweep_fail.log
for t in 64 128 256; do
LMPC_CLS3_THREADS=$t LMPC_DEBUG_STATS=1 python tools_profile.py 2048 reference_a1 2 stand 2>&1 | tail -2
done
