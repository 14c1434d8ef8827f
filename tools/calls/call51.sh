#!/bin/bash
# GPU call 51 (one B200): tests, smoke, launch lists, ncu counters + full captures of the final kernels
set -x
mkdir -p gpurun_out
( time timeout 900 python -m pytest tests -m gpu -x -q ) > gpurun_out/c51_pytest.log 2>&1
tail -4 gpurun_out/c51_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/c51_smoke.log 2>&1; tail -1 gpurun_out/c51_smoke.log
timeout 200 python tools/pass_shard.py --world 1 --rank 0 --once > gpurun_out/c51_once.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/c51_launches_w1.csv python tools/pass_shard.py --world 1 --rank 0 --once > /dev/null 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/c51_launches_w8.csv python tools/pass_shard.py --world 8 --rank 4 --once > /dev/null 2>&1
ncu --metrics smsp__inst_executed.sum,smsp__inst_executed_pipe_alu.sum,smsp__inst_executed_pipe_fma.sum,smsp__inst_executed_pipe_fp64.sum,smsp__inst_executed_pipe_xu.sum,smsp__inst_executed_pipe_lsu.sum,dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/c51_counters.csv python tools/pass_shard.py --world 1 --rank 0 --once > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k regex:'eval_inter_fused|eval_intra_warp|state_route_tables|position_unit_vectors' -o gpurun_out/c51_pass_full python tools/pass_shard.py --world 1 --rank 0 --once > gpurun_out/c51_ncu_full.log 2>&1
timeout 200 python tools/eta_profile.py > gpurun_out/c51_eta.log 2>&1 && \
ncu --set full --clock-control none -k regex:eta_score -c 2 -o gpurun_out/c51_eta_full python tools/eta_profile.py > gpurun_out/c51_ncu_eta.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/c51_launches_ls1000.csv python tools/ls_small.py --stops 1000 --moves 24 --repeat 1 > /dev/null 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/c51_launches_ls100.csv python tools/ls_small.py --stops 100 --vehicles 1 --moves 24 --repeat 1 > /dev/null 2>&1
ls -la gpurun_out | tail -20
