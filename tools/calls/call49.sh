#!/bin/bash
# tile kernel with no route test in tiles beyond the last row's route: all GPU tests, then A/B against HEAD (one B200)
set -x
mkdir -p gpurun_out
cp tools/calls/lib_b.so routest_b200/libroutest_cuda.so
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/c49_tests.log
cat gpurun_out/c49_tests.log
for round in 1 2; do for v in a b; do cp tools/calls/lib_$v.so routest_b200/libroutest_cuda.so; for w in 1 8; do echo "{\"variant\": \"$v\"}" >> gpurun_out/c49_ab.jsonl; timeout 200 python tools/pass_shard.py --world $w --rank 0 >> gpurun_out/c49_ab.jsonl 2>&1; done; done; done
cp tools/calls/lib_b.so routest_b200/libroutest_cuda.so
