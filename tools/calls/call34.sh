#!/bin/bash
# one-move search at small sizes with the merged apply + scan + tables launch (one B200); parity first
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/c34_tests.log
cat gpurun_out/c34_tests.log
for s in 10 100 400 1000 1500; do
  v=$((s/100)); [ $v -lt 1 ] && v=1
  for k in "-1 0 1" "-1 1 1" "4096 0 1" "4096 1 1"; do set -- $k
    timeout 120 python tools/ls_small.py --stops $s --vehicles $v --repeat 5 --state-fused-max $1 --apply-prep $2 --small-strips $3 >> gpurun_out/c34_ls.jsonl 2>&1
  done
done
timeout 120 python tools/ls_small.py --stops 1000 --vehicles 40 --repeat 5 --apply-prep 0 >> gpurun_out/c34_ls.jsonl 2>&1
timeout 120 python tools/ls_small.py --stops 1000 --vehicles 40 --repeat 5 --apply-prep 1 >> gpurun_out/c34_ls.jsonl 2>&1
cat gpurun_out/c34_ls.jsonl
timeout 200 python tools/request_latency.py > gpurun_out/c34_requests.log 2>&1; cat gpurun_out/c34_requests.log
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/c34_launches_ls1000.csv python tools/ls_small.py --stops 1000 --moves 24 --repeat 1 > /dev/null 2>&1
