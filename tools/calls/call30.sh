#!/bin/bash
# parity of the working-tree build, then A/B of two builds of the library in one box: a = previous, b = working tree
set -x
mkdir -p gpurun_out
cp tools/calls/lib_b.so routest_b200/libroutest_cuda.so
timeout 900 python -m pytest tests/test_moves_gpu.py tests/test_large_gpu.py -m gpu -x -q 2>&1 | tail -15 > gpurun_out/c30_tests.log
cat gpurun_out/c30_tests.log
for round in 1 2; do for v in a b; do cp tools/calls/lib_$v.so routest_b200/libroutest_cuda.so; for w in 1 8; do echo "{\"variant\": \"$v\"}" >> gpurun_out/c30_ab.jsonl; timeout 200 python tools/pass_shard.py --world $w --rank 0 >> gpurun_out/c30_ab.jsonl 2>&1; done; done; done
cp tools/calls/lib_b.so routest_b200/libroutest_cuda.so
cat gpurun_out/c30_ab.jsonl
