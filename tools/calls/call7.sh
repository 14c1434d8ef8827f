#!/bin/bash
# GPU call 7 (one B200): vote-gated updates, FAST cells, unroll 3
set -x
mkdir -p gpurun_out
( time timeout 1200 python -m pytest tests -m gpu -x -q ) > gpurun_out/c7_pytest.log 2>&1
tail -5 gpurun_out/c7_pytest.log
for w in 1 2 4 8; do timeout 200 python tools/pass_shard.py --fused 1 --world $w --rank $((w/2)) >> gpurun_out/c7_pass_shard.jsonl 2>&1; done
timeout 200 python tools/pass_shard.py --fused 0 --world 1 --rank 0 >> gpurun_out/c7_pass_shard.jsonl 2>&1
cat gpurun_out/c7_pass_shard.jsonl
timeout 200 python tools/pass_shard.py --world 1 --rank 0 --once > gpurun_out/c7_once.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'eval_inter_fused' -o gpurun_out/c7_fused_full python tools/pass_shard.py --world 1 --rank 0 --once > gpurun_out/c7_ncu3.log 2>&1
timeout 400 python bench.py --steps 20 --warmup 3 > gpurun_out/c7_bench_n1.json 2> gpurun_out/c7_bench_n1.err
head -c 4500 gpurun_out/c7_bench_n1.json; tail -3 gpurun_out/c7_bench_n1.err
ncu --metrics smsp__inst_executed.sum,smsp__inst_executed_pipe_alu.sum,smsp__inst_executed_pipe_fma.sum,smsp__inst_executed_pipe_fp64.sum,smsp__inst_executed_pipe_xu.sum,smsp__inst_executed_pipe_lsu.sum,dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/c7_counters.csv python tools/pass_shard.py --world 1 --rank 0 --once > gpurun_out/c7_ncu4.log 2>&1
