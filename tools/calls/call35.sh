#!/bin/bash
# one-move search: tables inside the merged launch where routes <= 32; the 1,500-stop step (one B200)
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_moves_gpu.py tests/test_large_gpu.py -m gpu -x -q 2>&1 | tail -3 > gpurun_out/c35_tests.log
cat gpurun_out/c35_tests.log
for s in 100 400 1000 1200 1500; do
  timeout 120 python tools/ls_small.py --stops $s --vehicles $((s/100)) --repeat 5 >> gpurun_out/c35_ls.jsonl 2>&1
done
cat gpurun_out/c35_ls.jsonl
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/c35_launches_ls1500.csv python tools/ls_small.py --stops 1500 --vehicles 15 --moves 24 --repeat 1 > /dev/null 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/c35_launches_ls1000.csv python tools/ls_small.py --stops 1000 --vehicles 10 --moves 24 --repeat 1 > /dev/null 2>&1
