"""torchrun check of the replicated multi-GPU local search (routest_b200/dist.py) against the single-GPU search.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 \
      tools/dist_search_check.py --stops 2000 --vehicles 20
Every rank builds the same instance, the ranks search it together (NCCL MIN of one int64 key per iteration), rank 0
also searches it alone; moves, cost and final routes must be identical.
"""
import argparse
import json
import os
import sys
import time
import zlib

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import DIST_SCALE, TS_CAR, synth_coords, synth_windows  # noqa: E402
from routest_b200 import Engine  # noqa: E402
from routest_b200.dist import DistributedSearch  # noqa: E402


def build(eng, seed, stops, vehicles):
    n = stops + 1
    lat, lon, demand, cap = synth_coords(seed, stops, vehicles)
    pr = eng.problem(n, 1)
    pr.set_coords(lat, lon)
    pr.build_matrix(DIST_SCALE, TS_CAR)
    _, t = pr.get_matrix(0)
    st = np.zeros((1, n, 4), np.int32)
    st[0, :, :3] = synth_windows(seed, t[0, 1:], t[1:, 0])
    st[0, :, 3] = demand
    pr.set_stops(st, [cap])
    trips, status = pr.greedy(demand[None, :].astype(float), [float(cap)], use_tw=True)
    assert status[0] == 0
    pr.set_routes(trips)
    return pr, trips


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--stops", type=int, default=2000)
    ap.add_argument("--vehicles", type=int, default=20)
    ap.add_argument("--max-moves", type=int, default=1 << 30)
    ap.add_argument("--exchange", default="nccl", choices=["nccl", "peer"])
    ap.add_argument("--batched", action="store_true", help="SPEC.md 8.1: a set of independent moves per iteration")
    a = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    eng = Engine(local)
    pr, trips = build(eng, 5678, a.stops, a.vehicles)
    search = DistributedSearch(pr, 1, device="cuda", check_every=8 if a.batched else 32, exchange=a.exchange,
                               batched=a.batched)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    moves, cost = search.run(a.max_moves)
    torch.cuda.synchronize()
    dt_multi = time.perf_counter() - t0
    routes_multi = pr.get_routes()[0]
    ok = True
    if world > 1:
        sig = torch.tensor([int(moves[0]), int(cost[0]), zlib.crc32(json.dumps(routes_multi).encode()) % (2 ** 31)], device="cuda")
        gathered = [torch.zeros_like(sig) for _ in range(world)]
        dist.all_gather(gathered, sig)
        ok = all(bool((g == gathered[0]).all()) for g in gathered)
    if rank == 0:
        pr1, _ = build(eng, 5678, a.stops, a.vehicles)
        t0 = time.perf_counter()
        m1, c1, _ = pr1.local_search(max_moves=a.max_moves, batched=a.batched)
        dt_single = time.perf_counter() - t0
        same = int(m1[0]) == int(moves[0]) and int(c1[0]) == int(cost[0]) and pr1.get_routes()[0] == routes_multi
        print(json.dumps({"world": world, "exchange": a.exchange, "batched": a.batched, "stops": a.stops, "moves": int(moves[0]), "cost": int(cost[0]),
                          "ranks_agree": ok, "equals_single_gpu": bool(same), "s_multi": round(dt_multi, 4),
                          "s_single": round(dt_single, 4), "us_per_move_multi": round(1e6 * dt_multi / max(1, int(moves[0])), 1),
                          "us_per_move_single": round(1e6 * dt_single / max(1, int(m1[0])), 1)}))
        assert ok and same
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
