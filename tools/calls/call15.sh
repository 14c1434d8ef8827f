#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 200 python tools/iter_shard.py --world 8 --rank 4 --once > gpurun_out/c15_once8.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/c15_launches_w8.csv python tools/iter_shard.py --world 8 --rank 4 --once > gpurun_out/c15_ncu.log 2>&1
