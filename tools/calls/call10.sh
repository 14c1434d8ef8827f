#!/bin/bash
# GPU call 10 (one B200): scan rides along with the vector kernel, walks as warp scans
set -x
mkdir -p gpurun_out
( time timeout 1200 python -m pytest tests -m gpu -x -q ) > gpurun_out/c10_pytest.log 2>&1
tail -5 gpurun_out/c10_pytest.log
for w in 1 2 4 8; do timeout 200 python tools/pass_shard.py --fused 1 --world $w --rank $((w/2)) >> gpurun_out/c10_pass_shard.jsonl 2>&1; done
cat gpurun_out/c10_pass_shard.jsonl
timeout 200 python tools/pass_shard.py --world 8 --rank 4 --once > gpurun_out/c10_once8.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/c10_launches_w8.csv python tools/pass_shard.py --world 8 --rank 4 --once > gpurun_out/c10_ncu2.log 2>&1
tail -8 gpurun_out/c10_launches_w8.csv | cut -d, -f5,9,15
