"""torchrun worker of tests/test_multi_gpu.py: one process per GPU, real NCCL and real CUDA IPC.

Every rank builds the same instance, the ranks search it together through routest_b200.dist.DistributedSearch, and
rank 0 checks the result against the CPU oracle (key for key) and against the single-GPU driver.  Prints one JSON line
per case; exits non-zero on the first mismatch.
"""
import json
import os
import sys
import zlib

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import TS_CAR, synth_vrptw  # noqa: E402
from oracle import oracle as O  # noqa: E402
from routest_b200 import ALL_FAMILIES, Engine  # noqa: E402
from routest_b200.dist import DistributedSearch, eta_score_distributed, eta_score_shard_dev  # noqa: E402


def problem(eng, inst, raw):
    pr = eng.problem(inst.n, 1)
    pr.set_coords(raw["lat"], raw["lon"])
    pr.build_matrix(10.0, TS_CAR)
    pr.set_stops(inst.stops[None], [inst.cap])
    pr.set_routes([raw["trips"]])
    return pr


def agree(values):
    t = torch.tensor(values, device="cuda", dtype=torch.int64)
    lo, hi = t.clone(), t.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN)
    dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    return bool((lo == hi).all().item())


def search_case(eng, rank, world, stops, vehicles, exchange, batched, max_iters, graph):
    inst, node, raw = synth_vrptw(5678, stops, vehicles)
    pr = problem(eng, inst, raw)
    search = DistributedSearch(pr, 1, device="cuda", check_every=4, exchange=exchange, batched=batched, graph=graph)
    moves, cost = search.run(max_iters)
    routes = pr.get_routes()[0]
    same = agree([int(moves[0]), int(cost[0]), zlib.crc32(json.dumps(routes).encode())])
    ok = same
    detail = {}
    if rank == 0:
        th = O.num_threads()
        if batched:
            final, okeys, sizes, ocost = O.local_search_batched(inst, node, max_iters=max_iters, threads=th)
            n_moves = int(okeys.size)
        else:
            final, okeys, ocost = O.local_search(inst, node, max_moves=max_iters, threads=th)
            n_moves = len(okeys)
        otrips, cur = [], []
        for c in final[1:]:            # same reading as rt_problem_get_routes: a separator closes a (possibly empty) route
            if c == 0:
                otrips.append(cur)
                cur = []
            else:
                cur.append(int(c))
        detail = {"oracle_moves": n_moves, "oracle_cost": int(ocost)}
        ok = ok and int(moves[0]) == n_moves and int(cost[0]) == int(ocost) and routes == otrips
    print(json.dumps({"case": "search", "world": world, "stops": stops, "exchange": exchange, "batched": batched,
                      "graph": bool(search._graph is not None), "moves": int(moves[0]), "cost": int(cost[0]),
                      "ranks_agree": same, "ok": bool(ok), **detail}), flush=True) if rank == 0 else None
    pr.close()
    return ok


def pass_case(eng, rank, world):
    """The bench's headline step at a size the oracle answers at once: plan_upload_dev on this rank's shard + NCCL MIN."""
    inst, node, raw = synth_vrptw(4567, 1500, 15)
    pr = problem(eng, inst, raw)
    pr.set_shard(rank, world)
    keys = torch.full((1,), 2 ** 63 - 1, dtype=torch.int64, device="cuda")
    s = torch.cuda.current_stream().cuda_stream
    nd = np.zeros((1, pr.pstride), np.int32)
    nd[0, :node.size] = node
    call = pr.plan_upload_call(np.ascontiguousarray(raw["lat"][None]), np.ascontiguousarray(raw["lon"][None]),
                               np.ascontiguousarray(inst.stops[None]), np.array([inst.cap], np.int32),
                               np.full(1, 2 ** 62, np.int64), nd, np.array([node.size], np.int32), keys.data_ptr(),
                               dist_scale=10.0, time_scale=TS_CAR, stream=s)
    call()
    dist.all_reduce(keys, op=dist.ReduceOp.MIN)
    got = (int(keys.item()) ^ -(2 ** 63)) & (2 ** 64 - 1)
    want = O.best_move(inst, node, threads=O.num_threads())
    ok = got == want
    if rank == 0:
        print(json.dumps({"case": "pass", "world": world, "key": got, "oracle": int(want), "ok": bool(ok)}), flush=True)
    pr.close()
    return ok


def eta_case(eng, rank, world):
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from test_eta import synth_model, synth_pairs
    _, blob = synth_model(seed=6, n_estimators=40, depth=5, n=2000)
    eng.eta_load_model(blob)
    cols = synth_pairs(99, 100_003)
    want = O.eta_eval(blob, O.eta_features(**cols), threads=O.num_threads())
    got = eta_score_distributed(eng, cols["distance_m"], cols["weather"], cols["traffic"], cols["weekday"], cols["hour"],
                                cols["age"]).numpy()
    ok = np.array_equal(got, want)
    # device pointers: every rank writes only its range of the shared-shape output
    dev = [torch.tensor(cols[k], device="cuda") for k in ("distance_m", "weather", "traffic", "weekday", "hour", "age")]
    out = torch.full((len(want),), -1.0, dtype=torch.float32, device="cuda")
    lo, hi = eta_score_shard_dev(eng, [d.data_ptr() for d in dev], len(want), out.data_ptr(), rank, world,
                                 torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    o = out.cpu().numpy()
    ok = ok and np.array_equal(o[lo:hi], want[lo:hi]) and (o[:lo] == -1).all() and (o[hi:] == -1).all()
    if rank == 0:
        print(json.dumps({"case": "eta", "world": world, "pairs": len(want), "ok": bool(ok)}), flush=True)
    return ok


def main():
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    eng = Engine(local)
    ok = True
    ok &= pass_case(eng, rank, world)
    for exchange, batched, graph in (("nccl", False, True), ("nccl", False, False), ("peer", False, True),
                                     ("peer", False, False), ("nccl", True, True), ("peer", True, True),
                                     ("peer", True, False)):
        ok &= search_case(eng, rank, world, 600, 8, exchange, batched, 24 if not batched else 6, graph)
    ok &= search_case(eng, rank, world, 5000, 50, "nccl", True, 4, True)    # multi-launch state tables, 140 routes
    ok &= search_case(eng, rank, world, 5000, 50, "peer", True, 4, True)
    ok &= eta_case(eng, rank, world)
    all_ok = agree([1 if ok else 0]) and ok
    eng.close()
    dist.destroy_process_group()
    sys.exit(0 if all_ok else 1)


if __name__ == "__main__":
    main()
