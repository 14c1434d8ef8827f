#!/bin/bash
# GPU call 52 (one B200): both bench arms and the secondary timings of the final kernels
set -x
mkdir -p gpurun_out
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/c52_bench_n1.json 2> gpurun_out/c52_bench_n1.err
head -c 900 gpurun_out/c52_bench_n1.json; echo; tail -2 gpurun_out/c52_bench_n1.err
timeout 300 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/c52_bench_ref.json 2>&1
head -c 600 gpurun_out/c52_bench_ref.json; echo
for w in 1 2 4 8; do for r in 0 $((w-1)); do timeout 200 python tools/pass_shard.py --world $w --rank $r >> gpurun_out/c52_pass_shard.jsonl 2>&1; done; done
timeout 200 python tools/pass_shard.py --fused 0 --world 1 --rank 0 >> gpurun_out/c52_pass_shard.jsonl 2>&1
for w in 1 2 8; do timeout 200 python tools/iter_shard.py --world $w --rank $((w/2)) >> gpurun_out/c52_iter.jsonl 2>&1; done
timeout 300 python tools/ls_bench.py > gpurun_out/c52_ls_bench.log 2>&1
timeout 200 python tools/request_latency.py > gpurun_out/c52_requests.log 2>&1
for cfg in "10 1 2 4" "100 1 20 70" "400 4 100 400" "1000 10 200 1000" "2000 20 200 1000" "4000 40 200 1000"; do set -- $cfg
  timeout 120 python tools/ls_small.py --stops $1 --vehicles $2 --repeat 5 --marginal $3 $4 >> gpurun_out/c52_ls_marginal.jsonl 2>&1
done
cat gpurun_out/c52_ls_marginal.jsonl gpurun_out/c52_ls_bench.log gpurun_out/c52_requests.log
