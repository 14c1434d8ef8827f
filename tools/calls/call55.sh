#!/bin/bash
# GPU call 55 (eight B200s): bench at N = 8, the 10,000-stop batched solve over 8 replicas
set -x
mkdir -p gpurun_out
run() { timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $1 "${@:2}"; }
run 29621 bench.py --gpus 8 --steps 20 --warmup 3 > gpurun_out/c55_bench_n8.json 2> gpurun_out/c55_bench_n8.err
head -c 300 gpurun_out/c55_bench_n8.json; echo; grep -v "^$" gpurun_out/c55_bench_n8.err | grep -i "error\|assert" | head -5
for ex in peer nccl; do
  run 29623 tools/dist_search_check.py --stops 10000 --vehicles 100 --batched --exchange $ex > gpurun_out/c55_solve10k_$ex.json 2>&1; tail -1 gpurun_out/c55_solve10k_$ex.json
done
