#!/bin/bash
# launch list of the one-move search at 1,000 and 100 stops (one B200)
set -x
mkdir -p gpurun_out
for s in 100 400 1000; do timeout 120 python tools/ls_small.py --stops $s --vehicles $((s/100)) >> gpurun_out/c32_ls.jsonl 2>&1; done
cat gpurun_out/c32_ls.jsonl
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/c32_launches_ls1000.csv python tools/ls_small.py --stops 1000 --moves 24 --repeat 1 > /dev/null 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/c32_launches_ls100.csv python tools/ls_small.py --stops 100 --vehicles 1 --moves 24 --repeat 1 > /dev/null 2>&1
