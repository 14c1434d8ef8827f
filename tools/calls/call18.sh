#!/bin/bash
# GPU call 18 (two B200s): full single-GPU suite on GPU 0, then the multi-GPU tests (peer table exchange included), then timings
set -x
mkdir -p gpurun_out
( time timeout 900 python -m pytest tests -m gpu -x -q ) > gpurun_out/c18_pytest.log 2>&1
tail -6 gpurun_out/c18_pytest.log
run() { timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $1 "${@:2}"; }
for ex in nccl peer; do
  run 29581 tools/dist_search_check.py --stops 5000 --vehicles 50 --batched --exchange $ex > gpurun_out/c18_solve5k_$ex.json 2>&1; tail -1 gpurun_out/c18_solve5k_$ex.json
  run 29582 tools/dist_search_check.py --stops 10000 --vehicles 100 --batched --exchange $ex > gpurun_out/c18_solve10k_$ex.json 2>&1; tail -1 gpurun_out/c18_solve10k_$ex.json
done
