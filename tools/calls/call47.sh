#!/bin/bash
# ncu full capture of the tile kernel (one B200)
set -x
mkdir -p gpurun_out
timeout 200 python tools/pass_shard.py --world 1 --rank 0 --once > gpurun_out/c47_once.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'eval_inter_fused' -o gpurun_out/c47_tile_full python tools/pass_shard.py --world 1 --rank 0 --once > gpurun_out/c47_ncu_full.log 2>&1
tail -3 gpurun_out/c47_ncu_full.log
