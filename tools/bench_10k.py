"""North-star configuration: ONE 10,000-stop VRPTW instance, matrix rows and move neighbourhood sharded across the
ranks, one NCCL MIN all-reduce of the best-move key per pass (BASELINE.json configs[3]).  Strong scaling.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29533 \
      tools/bench_10k.py --stops 10000 --steps 20
A pass = every rank rebuilds its rows of the route-order matrix (+ the replicated route state), scores its shard
of the full 2-opt/relocate/swap neighbourhood, and the ranks all-reduce one int64 key.  Timed with CUDA events on
the launching stream, max over ranks.  tests/perf/cpu_pass_10k.py times the CPU oracle on the same instance and prints
the same best key.
"""
import argparse
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import DIST_SCALE, TS_CAR, count_candidates, synth_coords, synth_windows  # noqa: E402
from routest_b200 import ALL_FAMILIES, Engine  # noqa: E402


def matrix_only(a, eng, rank, world, n, lat, lon):
    """Rank r builds rows [n*r/world, n*(r+1)/world) of the {dist,time} matrix into its own HBM (every rank holds the
    n unit vectors: 240 KB at 10,001 nodes).  Nothing is exchanged: a consumer that wants a full replica per GPU would
    all-gather 800 MB, the move kernels here rebuild their route-order rows instead."""
    r0, r1 = n * rank // world, n * (rank + 1) // world
    s = torch.cuda.current_stream().cuda_stream
    lat_d, lon_d = torch.tensor(lat, device="cuda"), torch.tensor(lon, device="cuda")
    xyz = torch.empty((3, n), dtype=torch.float64, device="cuda")
    cells = torch.empty((r1 - r0, eng.matrix_pitch(n), 2), dtype=torch.int32, device="cuda")

    def step():
        eng.unit_vectors_dev(lat_d.data_ptr(), lon_d.data_ptr(), n, 1, xyz.data_ptr(), s)
        eng.matrix_cells_dev(xyz.data_ptr(), n, 1, cells.data_ptr(), rows=(r0, r1), dist_scale=DIST_SCALE,
                             time_scale=TS_CAR, stream=s)

    for _ in range(5):
        step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
        torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1) / a.steps], device="cuda")
    check = cells[:, :n, 0].sum(dtype=torch.int64)      # a checksum of checksums: independent of the partition
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(check)
    if rank == 0:
        ms = float(ms.item())
        print(json.dumps({"workload": f"{n} x {n} distance/time matrix, rows sharded over {world} GPU(s), no collective",
                          "n_gpus": world, "ms": ms, "matrix_cells_per_s": n * n / ms * 1e3,
                          "GBs_aggregate": n * n * 8 / ms / 1e6, "sum_of_distances": int(check.item())}))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--stops", type=int, default=10000)
    ap.add_argument("--vehicles", type=int, default=100)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--matrix-only", action="store_true",
                    help="BASELINE config 4: only the node-order matrix, contiguous row blocks per rank, no collective")
    ap.add_argument("--exchange", default="nccl", choices=["nccl", "peer"],
                    help="nccl: all_reduce(MIN) per pass; peer: system-scope atomics into every replica's slot buffer")
    a = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    eng = Engine(local)
    n = a.stops + 1
    lat, lon, demand, cap = synth_coords(4567, a.stops, a.vehicles)
    if a.matrix_only:
        return matrix_only(a, eng, rank, world, n, lat, lon)
    pr = eng.problem(n, 1)
    pr.set_coords(lat, lon)
    pr.build_matrix(DIST_SCALE, TS_CAR)
    _, t = pr.get_matrix(0)
    st = np.zeros((1, n, 4), np.int32)
    st[0, :, :3] = synth_windows(4567, t[0, 1:], t[1:, 0])
    st[0, :, 3] = demand
    pr.set_stops(st, [cap])
    trips, status = pr.greedy(demand[None].astype(float), [float(cap)], use_tw=True)
    assert status[0] == 0
    pr.set_routes(trips)
    pr.set_shard(rank, world)
    P = n + len(trips[0])
    keys = torch.empty(1, dtype=torch.int64, device="cuda")
    s = torch.cuda.current_stream().cuda_stream

    if a.exchange == "peer":
        handle = pr.exchange_export()
        handles = [None] * world
        if world > 1:
            dist.all_gather_object(handles, handle)
        else:
            handles[0] = handle
        pr.exchange_connect(handles, rank, world)
        if world > 1:
            dist.barrier()

    def step():
        pr.refresh_dev(3, s)                                   # this rank's rows of the matrix + route state
        if a.exchange == "peer":
            pr.eval_moves_xchg_dev(ALL_FAMILIES, s)            # shard + MIN into every replica's slot over NVLink
            pr.collect_keys_xchg_dev(keys.data_ptr(), s)       # wait until all replicas have published
        else:
            pr.eval_moves_dev(keys.data_ptr(), ALL_FAMILIES, (rank, world), s)
            if world > 1:
                dist.all_reduce(keys, op=dist.ReduceOp.MIN)

    for _ in range(5):
        step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
        torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1) / a.steps], device="cuda")
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = float(ms.item())
    best = int(keys.item())
    if a.exchange == "peer":
        best ^= -(2 ** 63)  # the slot holds the unsigned packed key; fold it into the signed-order form used below
    if rank == 0:
        cand = count_candidates(trips[0], n)
        evals = sum(cand.values())
        line = {"exchange": a.exchange, "workload": f"one {a.stops}-stop VRPTW instance ({len(trips[0])} routes, P = {P}), rows and neighbourhood "
                            f"sharded over {world} GPU(s), one NCCL MIN all-reduce per pass",
                "n_gpus": world, "ms_per_pass": ms, "matrix_cells_per_s": P * P / ms * 1e3,
                "move_evals_per_s": evals / ms * 1e3, "cells_per_pass": P * P, "move_evals_per_pass": evals,
                "best_key": best & (2 ** 64 - 1)}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
