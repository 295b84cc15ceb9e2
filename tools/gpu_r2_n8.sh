# N-GPU lines (torchrun): headline (weak + strong + single-process sharded), config-4 sweep at 2^20 problems, config 5 at 1000 steps
N=${1:-8}
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512"
timeout 600 $TR bench.py --gpus $N --steps 10 --warmup 3 2> gpurun_out/n${N}_bench.err | grep '^{' > gpurun_out/n${N}_bench.json; cut -c1-200 gpurun_out/n${N}_bench.json
timeout 600 $TR bench.py --gpus $N --workload sweep --batch 131072 --steps 5 --warmup 3 2> gpurun_out/n${N}_sweep.err | grep '^{' > gpurun_out/n${N}_sweep.json; cut -c1-200 gpurun_out/n${N}_sweep.json
timeout 900 $TR bench.py --gpus $N --workload rollout --rollout-steps 1000 2> gpurun_out/n${N}_rollout.err | grep '^{' > gpurun_out/n${N}_rollout.json; cut -c1-300 gpurun_out/n${N}_rollout.json
