#!/bin/bash
# one-move search at small sizes: merged apply+vectors+scan launch, per-route tables instead of the one-block state
# kernel, short strips — each knob on and off (one B200); parity first
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_moves_gpu.py tests/test_large_gpu.py tests/test_greedy_gpu.py -m gpu -x -q 2>&1 | tail -15 > gpurun_out/c33_tests.log
cat gpurun_out/c33_tests.log
for s in 100 400 1000; do
  for k in "4096 0 0" "4096 0 1" "0 0 1" "4096 1 1" "0 1 1"; do set -- $k
    timeout 120 python tools/ls_small.py --stops $s --vehicles $((s/100)) --state-fused-max $1 --apply-prep $2 --small-strips $3 >> gpurun_out/c33_ls.jsonl 2>&1
  done
done
cat gpurun_out/c33_ls.jsonl
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/c33_launches_ls1000.csv python tools/ls_small.py --stops 1000 --moves 24 --repeat 1 --state-fused-max 0 > /dev/null 2>&1
