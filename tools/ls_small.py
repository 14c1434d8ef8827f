"""One-move local search on a small instance: time per applied move (development aid; --moves bounds the run so a
launch list of it stays short)."""
import argparse, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from routest_b200 import Engine
from tools.dist_search_check import build

ap = argparse.ArgumentParser()
ap.add_argument("--stops", type=int, default=1000)
ap.add_argument("--vehicles", type=int, default=10)
ap.add_argument("--moves", type=int, default=1 << 30)
ap.add_argument("--full", type=int, default=0)
ap.add_argument("--repeat", type=int, default=3)
ap.add_argument("--state-fused-max", type=int, default=-1)
ap.add_argument("--apply-prep", type=int, default=1)
ap.add_argument("--small-strips", type=int, default=1)
ap.add_argument("--persistent", type=int, default=1)
ap.add_argument("--persist-blocks", type=int, default=0)
ap.add_argument("--marginal", type=int, nargs=2, default=None, help="two move counts: time per move between them")
a = ap.parse_args()
eng = Engine(0)
eng.debug_set("state_fused_max", a.state_fused_max)
eng.debug_set("apply_prep", a.apply_prep)
eng.debug_set("small_strips", a.small_strips)
eng.debug_set("persistent_search", a.persistent)
eng.debug_set("persist_blocks", a.persist_blocks)


def timed(max_moves):
    best = None
    for _ in range(a.repeat):
        pr, trips = build(eng, 5678, a.stops, a.vehicles)
        t0 = time.perf_counter()
        moves, cost, _ = pr.local_search(max_moves=max_moves, full=bool(a.full))
        dt = time.perf_counter() - t0
        pr.close()
        if best is None or dt < best[0]:
            best = (dt, int(moves[0]), int(cost[0]))
    return best


if a.marginal:
    lo, hi = timed(a.marginal[0]), timed(a.marginal[1])
    print(json.dumps({"knobs": [a.persistent, a.persist_blocks, a.state_fused_max, a.apply_prep, a.small_strips],
                      "stops": a.stops, "vehicles": a.vehicles, "full": a.full, "moves": [lo[1], hi[1]],
                      "s": [round(lo[0], 5), round(hi[0], 5)],
                      "us_per_move_marginal": round(1e6 * (hi[0] - lo[0]) / max(1, hi[1] - lo[1]), 2)}))
    sys.exit(0)
best = None
for _ in range(a.repeat):
    pr, trips = build(eng, 5678, a.stops, a.vehicles)
    t0 = time.perf_counter()
    moves, cost, _ = pr.local_search(max_moves=a.moves, full=bool(a.full))
    dt = time.perf_counter() - t0
    pr.close()
    if best is None or dt < best[0]:
        best = (dt, int(moves[0]), int(cost[0]))
print(json.dumps({"knobs": [a.persistent, a.state_fused_max, a.apply_prep, a.small_strips], "stops": a.stops, "vehicles": a.vehicles, "full": a.full, "moves": best[1], "cost": best[2],
                  "s": round(best[0], 5), "us_per_move": round(1e6 * best[0] / max(1, best[1]), 2)}))
