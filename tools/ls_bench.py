"""Full local-search solve time: incremental (per-route-pair table) vs re-scoring everything vs batched iterations
(SPEC.md §8.1) — development aid."""
import json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from routest_b200 import Engine
from tools.dist_search_check import build

eng = Engine(0)
for stops, veh, cap_full in ((1000, 10, None), (5000, 50, None), (10000, 100, 512)):
    out = {"stops": stops, "vehicles": veh}
    for full in (False, True):
        pr, trips = build(eng, 5678, stops, veh)
        mm = (cap_full or 1 << 30) if full else 1 << 30
        t0 = time.perf_counter()
        moves, cost, _ = pr.local_search(max_moves=mm, full=full)
        dt = time.perf_counter() - t0
        out["full" if full else "incremental"] = {"moves": int(moves[0]), "cost": int(cost[0]), "s": round(dt, 4),
                                                  "us_per_move": round(1e6 * dt / max(1, int(moves[0])), 1)}
        pr.close()
    pr, trips = build(eng, 5678, stops, veh)
    t0 = time.perf_counter()
    moves, cost, _ = pr.local_search(batched=True)
    dt = time.perf_counter() - t0
    out["batched"] = {"moves": int(moves[0]), "cost": int(cost[0]), "s": round(dt, 4),
                      "us_per_move": round(1e6 * dt / max(1, int(moves[0])), 1)}
    pr.close()
    print(json.dumps(out))
