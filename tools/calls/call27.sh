#!/bin/bash
# GPU call 27 (eight B200s): NCCL small-message latency under different algorithm / protocol choices
set -x
mkdir -p gpurun_out
run() { timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $1 tools/nccl_latency.py; }
run 29601 2>/dev/null | grep world >> gpurun_out/c27_nccl.jsonl
NCCL_ALGO=Tree run 29602 2>/dev/null | grep world >> gpurun_out/c27_nccl.jsonl
NCCL_PROTO=LL run 29603 2>/dev/null | grep world >> gpurun_out/c27_nccl.jsonl
NCCL_ALGO=Tree NCCL_PROTO=LL run 29604 2>/dev/null | grep world >> gpurun_out/c27_nccl.jsonl
NCCL_NVLS_ENABLE=0 run 29605 2>/dev/null | grep world >> gpurun_out/c27_nccl.jsonl
NCCL_ALGO=Ring NCCL_PROTO=LL run 29606 2>/dev/null | grep world >> gpurun_out/c27_nccl.jsonl
cat gpurun_out/c27_nccl.jsonl
