# One box visit for the round's evidence: smoke, all GPU tests, bench lines of every workload, reference arm,
# ncu launch list of the bench command, full ncu captures of the three kernel families.
# Usage: bash tools/gpu_r2_final.sh <tag>   (files land in gpurun_out/<tag>_*)
T=${1:-r02}
set -x
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; tail -1 gpurun_out/${T}_smoke.log
timeout 1500 python -m pytest tests -m gpu -q -x -s > gpurun_out/${T}_pytest.log 2>&1; grep -E "passed|failed" gpurun_out/${T}_pytest.log; grep "parity-full" gpurun_out/${T}_pytest.log > gpurun_out/${T}_parity_fullsize.txt
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; cut -c1-200 gpurun_out/${T}_bench.json
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/${T}_reference_arm.json 2>/dev/null; cut -c1-200 gpurun_out/${T}_reference_arm.json
for wl in c3 ref_h30 sweep; do timeout 600 python bench.py --workload $wl --steps 5 --warmup 3 > gpurun_out/${T}_bench_$wl.json 2> gpurun_out/${T}_bench_$wl.err; cut -c1-160 gpurun_out/${T}_bench_$wl.json; done
timeout 600 python bench.py --workload c1 > gpurun_out/${T}_c1_latency.json 2>/dev/null; cut -c1-300 gpurun_out/${T}_c1_latency.json
timeout 900 python bench.py --workload rollout --rollout-steps 1000 > gpurun_out/${T}_rollout.json 2>/dev/null; cut -c1-300 gpurun_out/${T}_rollout.json
# launch list of the bench command (cold-cache, serialised: shares, not absolutes)
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${T}_launches.csv python bench.py --steps 2 --warmup 3 --no-warm > gpurun_out/${T}_launches.log 2>&1
# full captures; summarised ON the box (the reports with imported source are ~20 MB each; gpurun_out is capped at 64 MiB)
cap() {  # cap <mangled regex> <name> <kernel substring for the stall tool> <main file> <tools_profile args...>
  K=$1; N=$2; SUB=$3; MAIN=$4; shift 4
  bash tools/gpu_r2_prof.sh $K ${T}_ncu_$N "$@" | tail -2
  python tools/ncu_summary.py gpurun_out/${T}_ncu_$N.ncu-rep > gpurun_out/${T}_ncu_full_$N.csv
  python tools/ncu_stalls.py gpurun_out/${T}_ncu_$N.ncu-rep legged_mpc_control_b200/liblmpc.so "$SUB" $MAIN 50 30 > gpurun_out/${T}_ncu_stalls_$N.txt 2>&1
  ncu -i gpurun_out/${T}_ncu_$N.ncu-rep --page raw --csv > gpurun_out/${T}_ncu_raw_$N.csv 2>/dev/null
  rm -f gpurun_out/${T}_ncu_$N.ncu-rep
}
cap solve_warp solve_warp_kernel solve_warp_kernel lmpc_warp_solve.cuh 65536 northstar_a1 3
cap ILi24ELi8 solve_tile24_kernel solve_tile_kernelILi24ELi8 lmpc_tile_solve.cuh 4096 reference_a1 2 trot
cap ILi48ELi4 solve_tile48_kernel solve_tile_kernelILi48ELi4 lmpc_tile_solve.cuh 2048 reference_a1 2 stand
cap ILi15ELi2 solve_tile15_kernel solve_tile_kernelILi15ELi2 lmpc_tile_solve.cuh 16384 northstar_go1_h20 2
