# Full ncu captures only (the last block of gpu_r2_final.sh).  Usage: bash tools/gpu_r2_caps.sh <tag>
T=${1:-r02}
mkdir -p gpurun_out
cap() {
  K=$1; N=$2; SUB=$3; MAIN=$4; shift 4
  bash tools/gpu_r2_prof.sh $K ${T}_ncu_$N "$@" | tail -2
  python tools/ncu_summary.py gpurun_out/${T}_ncu_$N.ncu-rep > gpurun_out/${T}_ncu_full_$N.csv
  python tools/ncu_stalls.py gpurun_out/${T}_ncu_$N.ncu-rep legged_mpc_control_b200/liblmpc.so "$SUB" $MAIN 50 30 > gpurun_out/${T}_ncu_stalls_$N.txt 2>&1
  rm -f gpurun_out/${T}_ncu_$N.ncu-rep
}
cap solve_warp solve_warp_kernel solve_warp_kernel lmpc_warp_solve.cuh 65536 northstar_a1 3
cap ILi24ELi8 solve_tile24_kernel solve_tile_kernelILi24ELi8 lmpc_tile_solve.cuh 4096 reference_a1 2 trot
cap ILi48ELi4 solve_tile48_kernel solve_tile_kernelILi48ELi4 lmpc_tile_solve.cuh 2048 reference_a1 2 stand
cap ILi15ELi2 solve_tile15_kernel solve_tile_kernelILi15ELi2 lmpc_tile_solve.cuh 16384 northstar_go1_h20 2
