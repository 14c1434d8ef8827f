#!/bin/bash
set -x
mkdir -p gpurun_out
ROUTEST_DEBUG=1 timeout 200 python tools/ls_bench.py > gpurun_out/c22_ls_bench.log 2>&1; cat gpurun_out/c22_ls_bench.log
timeout 200 python tools/ls_profile.py 5000 batched > gpurun_out/c22_lsp.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/c22_launches_ls5k.csv python tools/ls_profile.py 5000 batched > /dev/null 2>&1
