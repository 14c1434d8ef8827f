#!/bin/bash
# GPU call 43 (two B200s): bench at N = 2
set -x
mkdir -p gpurun_out
run() { timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $1 "${@:2}"; }
run 29611 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/c43_bench_n2.json 2> gpurun_out/c43_bench_n2.err
head -c 400 gpurun_out/c43_bench_n2.json; echo; grep -v "^$" gpurun_out/c43_bench_n2.err | grep -i "error\|assert" | head -5
