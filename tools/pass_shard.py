"""What ONE rank of an N-GPU pass does, timed on one GPU (development aid): the headline instance of bench.py with
rt_problem_set_shard(rank, world), each kernel group alone and the whole graph-replayed pass without the exchange.

  python tools/pass_shard.py --world 8 --rank 3 [--stops 10000] [--once]
--once runs a single pass and exits (for ncu)."""
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
    ap.add_argument("--stops", type=int, default=10000)
    ap.add_argument("--vehicles", type=int, default=100)
    ap.add_argument("--world", type=int, default=8)
    ap.add_argument("--rank", type=int, default=3)
    ap.add_argument("--reps", type=int, default=50)
    ap.add_argument("--once", action="store_true")
    ap.add_argument("--fused", type=int, default=1, help="0: write the route-order matrix and read it back")
    ap.add_argument("--strip", type=int, default=0)
    a = ap.parse_args()
    eng = Engine(0)
    eng.debug_set("fused_eval", a.fused)
    eng.debug_set("strip_rows", a.strip)
    pr, w = bench.build_problem(eng, bench.SEED_10K, a.stops, a.vehicles)
    pr.set_shard(a.rank, a.world)
    keys = torch.zeros(1, dtype=torch.int64, device="cuda")
    s = torch.cuda.current_stream().cuda_stream
    one = lambda st: pr.plan_dev(keys.data_ptr(), dist_scale=bench.DIST_SCALE, time_scale=bench.TS_CAR,
                                 families=ALL_FAMILIES, stream=st)
    one(s)
    torch.cuda.synchronize()
    if a.once:
        one(s)
        torch.cuda.synchronize()
        return

    def timed(fn, reps=a.reps):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps * 1e3

    replay, graphed = bench.capture(torch, one)
    out = {"stops": a.stops, "shard": [a.rank, a.world], "fused": a.fused, "strip": a.strip, "graphed": graphed,
           "us_matrix": timed(lambda: pr.refresh_dev(4, s)), "us_state": timed(lambda: pr.refresh_dev(2, s)),
           "us_refresh": timed(lambda: pr.refresh_dev(3, s)),
           "us_eval": timed(lambda: pr.eval_moves_dev(keys.data_ptr(), ALL_FAMILIES, (a.rank, a.world), s)),
           "us_pass_eager": timed(lambda: one(s)), "us_pass_graph": timed(replay)}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
