#!/bin/bash
# the one-move search as one cooperative launch: parity, then time per move against the graph-replayed chain (one B200)
set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_moves_gpu.py -m gpu -x -q -k "launch_variants or one_launch" 2>&1 | tail -15 > gpurun_out/c38_tests_new.log
cat gpurun_out/c38_tests_new.log
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/c38_tests.log
cat gpurun_out/c38_tests.log
for s in 10 100 400 1000 2000 4000; do
  v=$((s/100)); [ $v -lt 1 ] && v=1
  for k in 0 1; do
    timeout 120 python tools/ls_small.py --stops $s --vehicles $v --repeat 5 --moves 2000 --full 1 --persistent $k >> gpurun_out/c38_ls.jsonl 2>&1
  done
done
cat gpurun_out/c38_ls.jsonl
