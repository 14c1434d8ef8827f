#!/bin/bash
# GPU call 28 (one B200): the round's evidence — tests, smoke, both bench arms, launch lists, ncu full + counters
set -x
mkdir -p gpurun_out
( time timeout 900 python -m pytest tests -m gpu -x -q ) > gpurun_out/c28_pytest.log 2>&1
tail -4 gpurun_out/c28_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/c28_smoke.log 2>&1; tail -1 gpurun_out/c28_smoke.log
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/c28_bench_n1.json 2> gpurun_out/c28_bench_n1.err
head -c 700 gpurun_out/c28_bench_n1.json; echo; tail -2 gpurun_out/c28_bench_n1.err
timeout 300 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/c28_bench_ref.json 2>&1
timeout 200 python tools/pass_shard.py --world 1 --rank 0 --once > gpurun_out/c28_once.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/c28_launches_w1.csv python tools/pass_shard.py --world 1 --rank 0 --once > /dev/null 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/c28_launches_w8.csv python tools/pass_shard.py --world 8 --rank 4 --once > /dev/null 2>&1
ncu --metrics smsp__inst_executed.sum,smsp__inst_executed_pipe_alu.sum,smsp__inst_executed_pipe_fma.sum,smsp__inst_executed_pipe_fp64.sum,smsp__inst_executed_pipe_xu.sum,smsp__inst_executed_pipe_lsu.sum,dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/c28_counters.csv python tools/pass_shard.py --world 1 --rank 0 --once > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k regex:'eval_inter_fused|eval_intra_warp|state_route_tables|position_unit_vectors' -o gpurun_out/c28_pass_full python tools/pass_shard.py --world 1 --rank 0 --once > gpurun_out/c28_ncu_full.log 2>&1
timeout 200 python tools/eta_profile.py > gpurun_out/c28_eta.log 2>&1 && \
ncu --set full --clock-control none -k regex:eta_score -c 2 -o gpurun_out/c28_eta_full python tools/eta_profile.py > gpurun_out/c28_ncu_eta.log 2>&1
for w in 1 2 4 8; do for r in 0 $((w-1)); do timeout 200 python tools/pass_shard.py --world $w --rank $r >> gpurun_out/c28_pass_shard.jsonl 2>&1; done; done
timeout 200 python tools/pass_shard.py --fused 0 --world 1 --rank 0 >> gpurun_out/c28_pass_shard.jsonl 2>&1
for w in 1 2 8; do timeout 200 python tools/iter_shard.py --world $w --rank $((w/2)) >> gpurun_out/c28_iter.jsonl 2>&1; done
timeout 200 python tools/ls_bench.py > gpurun_out/c28_ls_bench.log 2>&1
timeout 200 python tools/request_latency.py > gpurun_out/c28_requests.log 2>&1
ls -la gpurun_out | tail -20
