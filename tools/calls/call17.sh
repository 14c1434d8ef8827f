#!/bin/bash
# GPU call 17 (two B200s): can the NCCL all-reduce be captured into the iteration graph, and what does it save
set -x
mkdir -p gpurun_out
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29553 tools/coll_latency.py > gpurun_out/c17_coll2.json 2>&1; tail -2 gpurun_out/c17_coll2.json
