#!/bin/bash
# 2-GPU confirmation: bench.py under torchrun (weak scaling, blocking + overlapped weight hand-over)
set -x
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node ${1:-2} --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus ${1:-2} --steps 10 --warmup 3 --no-extras > gpurun_out/bench_n${1:-2}.json 2> gpurun_out/bench_n${1:-2}.err
echo rc=$?
tail -c 3000 gpurun_out/bench_n${1:-2}.json
tail -5 gpurun_out/bench_n${1:-2}.err
