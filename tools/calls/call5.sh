#!/bin/bash
# GPU call 5 (one B200): tests with the no-G kernels + A/B timings fused vs write-then-read
set -x
mkdir -p gpurun_out
( time timeout 1200 python -m pytest tests -m gpu -x -q ) > gpurun_out/c5_pytest.log 2>&1
tail -25 gpurun_out/c5_pytest.log
for f in 1 0; do for w in 1 2 4 8; do timeout 200 python tools/pass_shard.py --fused $f --world $w --rank $((w/2)) >> gpurun_out/c5_pass_shard.jsonl 2>&1; done; done
for h in 16 24 48 64; do timeout 200 python tools/pass_shard.py --fused 1 --world 8 --rank 4 --strip $h >> gpurun_out/c5_pass_shard.jsonl 2>&1; done
for h in 32 48; do timeout 200 python tools/pass_shard.py --fused 1 --world 1 --rank 0 --strip $h >> gpurun_out/c5_pass_shard.jsonl 2>&1; done
cat gpurun_out/c5_pass_shard.jsonl
timeout 200 python tools/pass_shard.py --world 1 --rank 0 --once > gpurun_out/c5_once.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'eval_inter_fused' -o gpurun_out/c5_fused_full python tools/pass_shard.py --world 1 --rank 0 --once > gpurun_out/c5_ncu3.log 2>&1
