#!/bin/bash
# GPU call 12 (one B200): where a batched iteration goes on a 1/N shard of the 5,000-stop instance
set -x
mkdir -p gpurun_out
for w in 1 2 8; do timeout 200 python tools/iter_shard.py --world $w --rank $((w/2)) >> gpurun_out/c12_iter.jsonl 2>&1; done
cat gpurun_out/c12_iter.jsonl
timeout 200 python tools/iter_shard.py --world 8 --rank 4 --once > gpurun_out/c12_once8.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/c12_launches_w8.csv python tools/iter_shard.py --world 8 --rank 4 --once > gpurun_out/c12_ncu.log 2>&1
timeout 200 python tools/iter_shard.py --world 1 --rank 0 --once > gpurun_out/c12_once1.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/c12_launches_w1.csv python tools/iter_shard.py --world 1 --rank 0 --once > gpurun_out/c12_ncu1.log 2>&1
