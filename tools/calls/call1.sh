#!/bin/bash
# GPU call 1 (one B200): tests, bench line, per-shard timings, ncu launch list + full capture of one pass
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv
( time timeout 1200 python -m pytest tests -m gpu -x -q ) > gpurun_out/c1_pytest.log 2>&1
tail -5 gpurun_out/c1_pytest.log
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/c1_bench_n1.json 2> gpurun_out/c1_bench_n1.err
tail -c 3000 gpurun_out/c1_bench_n1.json; tail -5 gpurun_out/c1_bench_n1.err
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/c1_bench_ref.json 2>&1
for w in 1 2 4 8; do timeout 200 python tools/pass_shard.py --world $w --rank $((w/2)) >> gpurun_out/c1_pass_shard.jsonl 2>&1; done
cat gpurun_out/c1_pass_shard.jsonl
timeout 200 python tools/pass_shard.py --world 1 --rank 0 --once > gpurun_out/c1_once.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/c1_launches_w1.csv python tools/pass_shard.py --world 1 --rank 0 --once > gpurun_out/c1_ncu1.log 2>&1
timeout 200 python tools/pass_shard.py --world 8 --rank 4 --once > gpurun_out/c1_once8.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/c1_launches_w8.csv python tools/pass_shard.py --world 8 --rank 4 --once > gpurun_out/c1_ncu2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'eval_inter|eval_intra|matrix_kernel' -o gpurun_out/c1_pass_full python tools/pass_shard.py --world 1 --rank 0 --once > gpurun_out/c1_ncu3.log 2>&1
ls -la gpurun_out
