"""Greedy construction time: engine (device matrix) vs the CPU oracle (development aid).
Lives under tests/ because it runs the CPU oracle as the checker and the baseline beside the engine."""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from bench import DIST_SCALE, TS_CAR, synth_coords, synth_windows
from routest_b200 import Engine
from oracle import oracle as O
eng = Engine(0)
for stops, veh, B in ((1000, 10, 64), (5000, 50, 1), (10000, 100, 1)):
    n = stops + 1
    raw = [synth_coords(2345 + b, stops, veh) for b in range(B)]
    lat = np.stack([r[0] for r in raw]); lon = np.stack([r[1] for r in raw])
    demand = np.stack([r[2] for r in raw]).astype(float); cap = np.array([r[3] for r in raw], float)
    pr = eng.problem(n, B); pr.set_coords(lat, lon); pr.build_matrix(DIST_SCALE, TS_CAR)
    st = np.zeros((B, n, 4), np.int32)
    mats = []
    for b in range(B):
        d, t = pr.get_matrix(b); mats.append((d, t) if b < 2 else None)
        st[b, :, :3] = synth_windows(2345 + b, t[0, 1:], t[1:, 0]); st[b, :, 3] = demand[b]
    pr.set_stops(st, cap.astype(np.int32))
    out = {"stops": stops, "batch": B}
    for tw in (False, True):
        pr.greedy(demand, cap, use_tw=tw)
        t0 = time.perf_counter(); trips, status = pr.greedy(demand, cap, use_tw=tw); dt = time.perf_counter() - t0
        d, t = mats[0]
        t1 = time.perf_counter(); rc, otr = O.greedy_trips(d, demand[0], cap[0], 9e12, T=t if tw else None, tw=st[0, :, :3] if tw else None); dc = time.perf_counter() - t1
        out["tw" if tw else "plain"] = {"engine_ms_batch": round(dt * 1e3, 2), "oracle_ms_one_instance": round(dc * 1e3, 2),
                                        "same": otr == trips[0], "trips": len(otr)}
    print(json.dumps(out)); pr.close()
