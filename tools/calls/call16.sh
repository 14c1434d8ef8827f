#!/bin/bash
# GPU call 16 (eight B200s): bench at N=8 and N=4 (engine arm), collective latencies, the 10,000-stop batched solve over 8 ranks
set -x
mkdir -p gpurun_out
nvidia-smi -L | head -8
run() { timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port $2 "${@:3}"; }
run 8 29571 bench.py --gpus 8 --steps 20 --warmup 3 > gpurun_out/c16_bench_n8.json 2> gpurun_out/c16_bench_n8.err
head -c 1500 gpurun_out/c16_bench_n8.json; echo; tail -4 gpurun_out/c16_bench_n8.err
run 4 29572 bench.py --gpus 4 --steps 20 --warmup 3 --no-extra > gpurun_out/c16_bench_n4.json 2> gpurun_out/c16_bench_n4.err
head -c 600 gpurun_out/c16_bench_n4.json; echo
run 8 29573 tools/coll_latency.py > gpurun_out/c16_coll8.json 2>&1; tail -1 gpurun_out/c16_coll8.json
run 8 29574 tools/dist_search_check.py --stops 10000 --vehicles 100 --batched > gpurun_out/c16_solve10k_8.json 2>&1; tail -1 gpurun_out/c16_solve10k_8.json
run 8 29575 tools/dist_search_check.py --stops 5000 --vehicles 50 --batched > gpurun_out/c16_solve5k_8.json 2>&1; tail -1 gpurun_out/c16_solve5k_8.json
