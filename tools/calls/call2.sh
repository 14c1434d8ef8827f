#!/bin/bash
# GPU call 2 (one B200): tests after the state/intra/strip changes + per-shard timings + launch list of a 1/8 shard
set -x
mkdir -p gpurun_out
( time timeout 1200 python -m pytest tests -m gpu -x -q ) > gpurun_out/c2_pytest.log 2>&1
tail -5 gpurun_out/c2_pytest.log
for w in 1 2 4 8; do timeout 200 python tools/pass_shard.py --world $w --rank $((w/2)) >> gpurun_out/c2_pass_shard.jsonl 2>&1; done
cat gpurun_out/c2_pass_shard.jsonl
timeout 200 python tools/pass_shard.py --world 8 --rank 4 --once > gpurun_out/c2_once8.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/c2_launches_w8.csv python tools/pass_shard.py --world 8 --rank 4 --once > gpurun_out/c2_ncu2.log 2>&1
timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu > gpurun_out/c2_bench_n1.json 2> gpurun_out/c2_bench_n1.err
tail -c 1500 gpurun_out/c2_bench_n1.json; tail -5 gpurun_out/c2_bench_n1.err
