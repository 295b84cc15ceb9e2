#!/bin/bash
# DRAM traffic of the warp solve kernel with and without an L2 persistence window.
mkdir -p gpurun_out
M="--metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_bytes.sum"
python tools_profile.py 32768 northstar_a1 3 > gpurun_out/plainp.log 2>&1 &&
ncu $M --clock-control none -k regex:solve_warp -s 2 -c 1 --csv --log-file gpurun_out/traffic_a.csv python tools_profile.py 32768 northstar_a1 3 > /dev/null 2>&1
LMPC_L2_PERSIST=1 ncu $M --clock-control none -k regex:solve_warp -s 2 -c 1 --csv --log-file gpurun_out/traffic_b.csv python tools_profile.py 32768 northstar_a1 3 2>&1 | grep lmpc
LMPC_L2_PERSIST=1 python tools_profile.py 32768 northstar_a1 3 2>&1 | tail -1
grep -h -E "dram|duration|lts" gpurun_out/traffic_a.csv gpurun_out/traffic_b.csv | cut -d, -f13-
