# ncu full capture of one kernel: bash tools/gpu_r2_prof.sh <mangled-name regex> <out name> <tools_profile args...>
# (-s 2: the third matching launch — every solve call launches a kernel twice, first pass and the usually empty retry pass)
K=$1; OUT=$2; shift 2
mkdir -p gpurun_out
python tools_profile.py "$@" > gpurun_out/${OUT}_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on --kernel-name-base mangled -k regex:$K -s 2 -c 1 -f -o gpurun_out/$OUT python tools_profile.py "$@" > gpurun_out/${OUT}_ncu.log 2>&1
cat gpurun_out/${OUT}_plain.log; tail -3 gpurun_out/${OUT}_ncu.log
