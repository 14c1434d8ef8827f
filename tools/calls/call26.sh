#!/bin/bash
# GPU call 26 (one B200): bench line with the triangular kernel + where the e2e call's host time goes
set -x
mkdir -p gpurun_out
ROUTEST_DEBUG_PLAN=1 timeout 600 python bench.py --steps 100 --warmup 3 --no-extra --no-cpu > gpurun_out/c26_bench_dbg.json 2> gpurun_out/c26_bench_dbg.err
grep routest gpurun_out/c26_bench_dbg.err | tail -3
python - <<'PY'
import json
d=json.loads(open('gpurun_out/c26_bench_dbg.json').read().strip().splitlines()[-1])
print("ms", d["ms_per_step"], "e2e", d["e2e"]["ms_per_step"], "own", d["kernels"]["whole_pass_without_exchange"]["ms"], "wtr", d["kernels"]["whole_pass_write_then_read"]["ms"])
PY
