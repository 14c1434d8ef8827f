#!/bin/bash
# GPU call 3 (two B200s): the multi-GPU tests under real NCCL / IPC, bench at N=2 (both arms), strip-height sweep on GPU 0
set -x
mkdir -p gpurun_out
nvidia-smi -L
( time timeout 900 python -m pytest tests/test_multi_gpu.py -m gpu -x -q -s ) > gpurun_out/c3_pytest_mgpu.log 2>&1
tail -15 gpurun_out/c3_pytest_mgpu.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29551 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/c3_bench_n2.json 2> gpurun_out/c3_bench_n2.err
tail -c 6000 gpurun_out/c3_bench_n2.json; tail -8 gpurun_out/c3_bench_n2.err
for h in 16 24 32 48; do ROUTEST_STRIP_ROWS=$h timeout 200 python tools/pass_shard.py --world 8 --rank 4 >> gpurun_out/c3_strip_sweep.jsonl 2>&1; done
cat gpurun_out/c3_strip_sweep.jsonl
