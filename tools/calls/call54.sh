#!/bin/bash
# GPU call 54 (four B200s): bench at N = 4
set -x
mkdir -p gpurun_out
run() { timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port $1 "${@:2}"; }
run 29631 bench.py --gpus 4 --steps 20 --warmup 3 > gpurun_out/c54_bench_n4.json 2> gpurun_out/c54_bench_n4.err
head -c 400 gpurun_out/c54_bench_n4.json; echo; grep -v "^$" gpurun_out/c54_bench_n4.err | grep -i "error\|assert" | head -5
