#!/bin/bash
# one-move search: re-scoring everything (full) against the incremental table, 1,500 to 7,000 stops (one B200)
set -x
mkdir -p gpurun_out
for s in 1500 2000 3000 5000 7000; do for f in 0 1; do
  timeout 120 python tools/ls_small.py --stops $s --vehicles $((s/100)) --repeat 2 --moves 800 --full $f >> gpurun_out/c36_ls.jsonl 2>&1
done; done
cat gpurun_out/c36_ls.jsonl
