#!/bin/bash
# GPU call 6 (one B200): row-slot version of the no-G tile kernel
set -x
mkdir -p gpurun_out
( time timeout 1200 python -m pytest tests -m gpu -x -q ) > gpurun_out/c6_pytest.log 2>&1
tail -5 gpurun_out/c6_pytest.log
for w in 1 2 4 8; do timeout 200 python tools/pass_shard.py --fused 1 --world $w --rank $((w/2)) >> gpurun_out/c6_pass_shard.jsonl 2>&1; done
cat gpurun_out/c6_pass_shard.jsonl
timeout 200 python tools/pass_shard.py --world 1 --rank 0 --once > gpurun_out/c6_once.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'eval_inter_fused' -o gpurun_out/c6_fused_full python tools/pass_shard.py --world 1 --rank 0 --once > gpurun_out/c6_ncu3.log 2>&1
timeout 400 python bench.py --steps 20 --warmup 3 > gpurun_out/c6_bench_n1.json 2> gpurun_out/c6_bench_n1.err
head -c 2500 gpurun_out/c6_bench_n1.json; tail -3 gpurun_out/c6_bench_n1.err
