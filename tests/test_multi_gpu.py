"""GPU, two or more B200s: the sharded path under a REAL collective — NCCL MIN of the move keys, the CUDA-IPC peer
exchange (exchange.cu), the graph-replayed loop, batched iterations with the table all-reduce, the bench's headline
step (rt_problem_plan_upload_dev + all_reduce) and sharded ETA scoring — each against the CPU oracle.  One process per
GPU through torchrun (tests/mgpu/worker.py); skipped where fewer than two GPUs are visible."""
import json
import os
import subprocess
import sys

import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu


@pytest.mark.timeout(900)
def test_two_ranks_real_nccl_and_ipc_match_the_oracle():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (run with gpurun --gpus 2)")
    env = dict(os.environ, OMP_NUM_THREADS="8")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29541", os.path.join(ROOT, "tests", "mgpu", "worker.py")]
    res = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=850)
    lines = [json.loads(ln) for ln in res.stdout.splitlines() if ln.startswith("{")]
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    assert len(lines) == 11 and all(ln["ok"] for ln in lines), lines
    assert any(ln.get("graph") for ln in lines if ln["case"] == "search"), "no search ran from a CUDA graph"
