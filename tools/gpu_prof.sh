#!/bin/bash
# ncu full capture of the warp kernel on a small fixed workload (tools_profile.py).
mkdir -p gpurun_out
python tools_profile.py 32768 northstar_a1 3 > gpurun_out/plainp.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:solve_warp -s 2 -c 1 -f -o gpurun_out/prof_warp python tools_profile.py 32768 northstar_a1 3 > gpurun_out/ncu_full.log 2>&1
cat gpurun_out/plainp.log; tail -3 gpurun_out/ncu_full.log
