#!/bin/bash
# GPU call 8 (two B200s): multi-GPU tests, bench at N=2 (both arms), where an iteration of the replicated search goes
set -x
mkdir -p gpurun_out
( time timeout 900 python -m pytest tests/test_multi_gpu.py -m gpu -x -q ) > gpurun_out/c8_pytest_mgpu.log 2>&1
tail -5 gpurun_out/c8_pytest_mgpu.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29551 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/c8_bench_n2.json 2> gpurun_out/c8_bench_n2.err
head -c 3000 gpurun_out/c8_bench_n2.json; echo; tail -c 3500 gpurun_out/c8_bench_n2.json; tail -4 gpurun_out/c8_bench_n2.err
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29552 bench.py --impl reference --gpus 2 --steps 3 --warmup 1 > gpurun_out/c8_bench_ref_n2.json 2> gpurun_out/c8_bench_ref_n2.err
cut -c1-600 gpurun_out/c8_bench_ref_n2.json
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29553 tools/coll_latency.py > gpurun_out/c8_coll2.json 2>&1; tail -2 gpurun_out/c8_coll2.json
