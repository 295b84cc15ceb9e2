#!/bin/bash
mkdir -p gpurun_out
python tools_profile.py 8192 northstar_go1_h20 2 > gpurun_out/plain_t15.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:solve_tile -s 2 -c 1 -f -o gpurun_out/prof_t15 python tools_profile.py 8192 northstar_go1_h20 2 > gpurun_out/ncu_t15.log 2>&1
python tools_profile.py 2048 reference_a1 2 > gpurun_out/plain_t24.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:solve_tile -s 3 -c 1 -f -o gpurun_out/prof_t24 python tools_profile.py 2048 reference_a1 2 > gpurun_out/ncu_t24.log 2>&1
tail -1 gpurun_out/plain_t15.log gpurun_out/plain_t24.log; tail -2 gpurun_out/ncu_t15.log gpurun_out/ncu_t24.log
