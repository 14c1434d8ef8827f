"""What ONE rank of an N-GPU batched iteration (SPEC.md §8.1) does, timed on one GPU (development aid): the 5,000-stop
instance of bench.py with rt_problem_set_shard(rank, world); the table evaluation, the selection + apply + refresh, and
the graph-replayed pair.  --once runs three iterations and exits (for ncu)."""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from routest_b200 import ALL_FAMILIES, Engine  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--stops", type=int, default=5000)
    ap.add_argument("--vehicles", type=int, default=50)
    ap.add_argument("--world", type=int, default=8)
    ap.add_argument("--rank", type=int, default=3)
    ap.add_argument("--once", action="store_true")
    a = ap.parse_args()
    eng = Engine(0)
    pr, w = bench.build_problem(eng, bench.SEED_5K, a.stops, a.vehicles)
    pr.set_shard(a.rank, a.world)
    table = torch.empty(pr.table_entries(), dtype=torch.int64, device="cuda")
    s = torch.cuda.current_stream().cuda_stream
    ev = lambda st: pr.eval_table_dev(table.data_ptr(), ALL_FAMILIES, (a.rank, a.world), st)
    ap_ = lambda st: pr.apply_batch_dev(table.data_ptr(), st)
    ev(s)
    torch.cuda.synchronize()
    if a.once:
        for _ in range(3):   # a shard's table alone is not the full minimum, but the kernels and their sizes are the same
            ap_(s)
            ev(s)
        torch.cuda.synchronize()
        return

    def timed(fn, reps=30):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps * 1e3

    out = {"stops": a.stops, "shard": [a.rank, a.world], "us_eval_table": timed(lambda: ev(s))}
    both = lambda st: (ap_(st), ev(st))
    both(s)
    replay, graphed = bench.capture(torch, both)
    out["us_apply_refresh_eval_graph"] = timed(replay)   # early iterations: every batch is large
    out["graphed"] = graphed
    print(json.dumps(out))


if __name__ == "__main__":
    main()
