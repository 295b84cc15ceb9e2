# 2-GPU checks: torchrun bench line (weak + strong + sharded single-process), C++ sharded test on 2 devices
set -x
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/n2_bench.json 2> gpurun_out/n2_bench.err; echo "rc $?"; tail -3 gpurun_out/n2_bench.err; cut -c1-300 gpurun_out/n2_bench.json
timeout 600 python -m pytest tests/test_cpp_wrapper.py -m gpu -q -x 2>&1 | tail -3
