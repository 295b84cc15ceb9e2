#!/bin/bash
# final-code evidence: launch list + full capture of solve_warp_kernel (headline), full captures of the tile kernels
set -x
mkdir -p gpurun_out
python bench.py --steps 3 --warmup 3 --no-warm > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 3 --warmup 3 --no-warm > gpurun_out/ncu_launch.log 2>&1
python bench.py --steps 3 --warmup 3 --no-warm > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:solve_warp -s 3 -c 1 -f -o gpurun_out/prof_warp python bench.py --steps 3 --warmup 3 --no-warm > gpurun_out/ncu_full.log 2>&1
bash tools/gpu_prof_tile.sh
ls -la gpurun_out
