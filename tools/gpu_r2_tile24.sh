# Full ncu capture + stall-by-region summary of the class-24 kernel of the final binary (8 warps per QP).
# Usage: bash tools/gpu_r2_tile24.sh <tag>
T=${1:-r02}
mkdir -p gpurun_out
N=solve_tile24_kernel
bash tools/gpu_r2_prof.sh ILi24ELi8 ${T}_ncu_$N 4096 reference_a1 2 trot | tail -2
python tools/ncu_summary.py gpurun_out/${T}_ncu_$N.ncu-rep > gpurun_out/${T}_ncu_full_$N.csv
python tools/ncu_stalls.py gpurun_out/${T}_ncu_$N.ncu-rep legged_mpc_control_b200/liblmpc.so solve_tile_kernelILi24ELi8 lmpc_tile_solve.cuh 50 30 > gpurun_out/${T}_ncu_stalls_$N.txt 2>&1
rm -f gpurun_out/${T}_ncu_$N.ncu-rep
head -30 gpurun_out/${T}_ncu_stalls_$N.txt | cut -c1-180
