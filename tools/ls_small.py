"""Sequential local search on small instances: incremental vs full re-scoring driver (development aid)."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from routest_b200 import Engine
from tools.dist_search_check import build

eng = Engine(0)
for stops, veh in ((100, 10), (200, 10), (400, 10), (1000, 10), (2000, 20)):
    out = {"stops": stops}
    for name, kw in (("incremental", {}), ("full", {"full": True}), ("batched", {"batched": True})):
        best = 1e9
        for rep in range(3):
            pr, trips = build(eng, 5678, stops, veh)
            t0 = time.perf_counter()
            moves, cost, _ = pr.local_search(**kw)
            best = min(best, time.perf_counter() - t0)
            pr.close()
        out[name] = {"moves": int(moves[0]), "ms": round(best * 1e3, 2), "us_per_move": round(1e6 * best / max(1, int(moves[0])), 1)}
    print(json.dumps(out))
