#!/bin/bash
# GPU call 4 (one B200): tests with the warp-per-anchor kernel + shard timings
set -x
mkdir -p gpurun_out
( time timeout 1200 python -m pytest tests -m gpu -x -q ) > gpurun_out/c4_pytest.log 2>&1
tail -25 gpurun_out/c4_pytest.log
for w in 1 2 8; do timeout 200 python tools/pass_shard.py --world $w --rank $((w/2)) >> gpurun_out/c4_pass_shard.jsonl 2>&1; done
cat gpurun_out/c4_pass_shard.jsonl
timeout 200 python tools/pass_shard.py --world 8 --rank 4 --once > gpurun_out/c4_once8.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/c4_launches_w8.csv python tools/pass_shard.py --world 8 --rank 4 --once > gpurun_out/c4_ncu2.log 2>&1
timeout 200 python tools/coll_latency.py > gpurun_out/c4_coll1.json 2>&1; cat gpurun_out/c4_coll1.json
