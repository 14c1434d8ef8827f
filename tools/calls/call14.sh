#!/bin/bash
# GPU call 14 (one B200): fused selection with the table in shared memory up to 152 routes, signed table without a flip pass
set -x
mkdir -p gpurun_out
( time timeout 1200 python -m pytest tests -m gpu -x -q ) > gpurun_out/c14_pytest.log 2>&1
tail -5 gpurun_out/c14_pytest.log
for w in 1 2 8; do timeout 200 python tools/iter_shard.py --world $w --rank $((w/2)) >> gpurun_out/c14_iter.jsonl 2>&1; done
cat gpurun_out/c14_iter.jsonl
