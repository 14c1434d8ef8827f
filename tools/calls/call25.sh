#!/bin/bash
# GPU call 25 (one B200): work-balanced shard boundaries — every rank of 8 emulated on one GPU
set -x
mkdir -p gpurun_out
( time timeout 900 python -m pytest tests -m gpu -x -q ) > gpurun_out/c25_pytest.log 2>&1
tail -4 gpurun_out/c25_pytest.log
for r in 0 1 3 5 7; do timeout 200 python tools/pass_shard.py --world 8 --rank $r >> gpurun_out/c25_pass_shard.jsonl 2>&1; done
for r in 0 1; do timeout 200 python tools/pass_shard.py --world 2 --rank $r >> gpurun_out/c25_pass_shard.jsonl 2>&1; done
timeout 200 python tools/pass_shard.py --world 1 --rank 0 >> gpurun_out/c25_pass_shard.jsonl 2>&1
cat gpurun_out/c25_pass_shard.jsonl
