#!/bin/bash
# one-launch search with the last block applying the move: parity, then marginal time per move (one B200)
set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_moves_gpu.py -m gpu -x -q -k "launch_variants or one_launch" 2>&1 | tail -5 > gpurun_out/c39_tests_new.log
cat gpurun_out/c39_tests_new.log
for cfg in "100 1 20 70" "400 4 100 400" "1000 10 200 1000" "2000 20 200 1000"; do set -- $cfg
  for k in "0 0" "1 0" "1 1" "1 2"; do set -- $cfg $k
    timeout 120 python tools/ls_small.py --stops $1 --vehicles $2 --repeat 5 --full 1 --marginal $3 $4 --persistent $5 --persist-blocks $6 >> gpurun_out/c39_ls.jsonl 2>&1
  done
done
cat gpurun_out/c39_ls.jsonl
