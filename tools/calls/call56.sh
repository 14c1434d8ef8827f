#!/bin/bash
# GPU call 56 (two B200s): bench at N = 2 after the 5,000-stop solve block runs twice
set -x
mkdir -p gpurun_out
run() { timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $1 "${@:2}"; }
run 29611 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/c56_bench_n2.json 2> gpurun_out/c56_bench_n2.err
head -c 200 gpurun_out/c56_bench_n2.json; echo; grep -v "^$" gpurun_out/c56_bench_n2.err | grep -i "error\|assert" | head -5
