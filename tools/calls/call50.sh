#!/bin/bash
# A/B of two builds of the library (one B200): a = HEAD, b = working tree
set -x
mkdir -p gpurun_out
for round in 1 2; do for v in a b; do cp tools/calls/lib_$v.so routest_b200/libroutest_cuda.so; for w in 1 8; do echo "{\"variant\": \"$v\"}" >> gpurun_out/c50_ab.jsonl; timeout 200 python tools/pass_shard.py --world $w --rank 0 >> gpurun_out/c50_ab.jsonl 2>&1; done; done; done
cp tools/calls/lib_b.so routest_b200/libroutest_cuda.so
