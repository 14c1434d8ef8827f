#!/bin/bash
# GPU call 53 (two B200s): multi-GPU tests, bench at N = 2
set -x
mkdir -p gpurun_out
( time timeout 900 python -m pytest tests/test_multi_gpu.py -m gpu -x -q ) > gpurun_out/c53_pytest_mgpu.log 2>&1
tail -6 gpurun_out/c53_pytest_mgpu.log
run() { timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $1 "${@:2}"; }
run 29611 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/c53_bench_n2.json 2> gpurun_out/c53_bench_n2.err
head -c 300 gpurun_out/c53_bench_n2.json; echo; grep -v "^$" gpurun_out/c53_bench_n2.err | grep -i "error\|assert" | head -5
for ex in nccl peer; do
  run 29613 tools/dist_search_check.py --stops 10000 --vehicles 100 --batched --exchange $ex > gpurun_out/c53_solve10k_$ex.json 2>&1; tail -1 gpurun_out/c53_solve10k_$ex.json
done
