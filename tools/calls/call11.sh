#!/bin/bash
# GPU call 11 (one B200): two columns per thread in the no-matrix tile kernel
set -x
mkdir -p gpurun_out
( time timeout 1200 python -m pytest tests -m gpu -x -q ) > gpurun_out/c11_pytest.log 2>&1
tail -5 gpurun_out/c11_pytest.log
for c in 2 1; do for w in 1 2 4 8; do timeout 200 python tools/pass_shard.py --cols $c --world $w --rank $((w/2)) >> gpurun_out/c11_pass_shard.jsonl 2>&1; done; done
for h in 32 48; do timeout 200 python tools/pass_shard.py --cols 2 --world 1 --rank 0 --strip $h >> gpurun_out/c11_pass_shard.jsonl 2>&1; done
for h in 16 24 48; do timeout 200 python tools/pass_shard.py --cols 2 --world 8 --rank 4 --strip $h >> gpurun_out/c11_pass_shard.jsonl 2>&1; done
cat gpurun_out/c11_pass_shard.jsonl
timeout 200 python tools/pass_shard.py --world 8 --rank 4 --once > gpurun_out/c11_once8.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/c11_launches_w8.csv python tools/pass_shard.py --world 8 --rank 4 --once > gpurun_out/c11_ncu2.log 2>&1
