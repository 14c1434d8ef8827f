"""CPU side of tools/bench_10k.py: the same 10,000-stop instance (same seed, same greedy start) scored by the CPU
oracle with all host threads.  Prints the pass time and the best key, which must equal bench_10k.py's `best_key`.
Lives under tests/ because it runs the oracle; it needs no GPU (the oracle's greedy start is bit-identical to the
engine's, tests/test_greedy_gpu.py).

  python tests/perf/cpu_pass_10k.py --stops 10000
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from bench import DIST_SCALE, TS_CAR, count_candidates, synth_coords, synth_windows  # noqa: E402
from oracle import oracle as O  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--stops", type=int, default=10000)
    ap.add_argument("--vehicles", type=int, default=100)
    a = ap.parse_args()
    n = a.stops + 1
    lat, lon, demand, cap = synth_coords(4567, a.stops, a.vehicles)
    th = O.num_threads()
    t0 = time.perf_counter()
    m = O.matrix_rows(O.unit_vectors(lat, lon), 0, n, DIST_SCALE, TS_CAR, threads=th)
    t1 = time.perf_counter()
    st = np.zeros((n, 4), np.int32)
    st[:, :3] = synth_windows(4567, m["time"][0, 1:], m["time"][1:, 0])
    st[:, 3] = demand
    rc, trips = O.greedy_trips(m["dist"], demand.astype(float), float(cap), 9e12, T=m["time"], tw=st[:, :3])
    assert rc == 0
    inst = O.Instance(m["dist"], m["time"], st, cap)
    node = O.build_positions(trips, n)
    t2 = time.perf_counter()
    k = O.best_move(inst, node, threads=th)
    t3 = time.perf_counter()
    evals = sum(count_candidates(trips, n).values())
    P = node.size
    print(json.dumps({"stops": a.stops, "routes": len(trips), "P": int(P), "threads": th, "matrix_s": t1 - t0,
                      "best_move_s": t3 - t2, "pass_s": (t1 - t0) + (t3 - t2),
                      "matrix_cells_per_s": n * n / (t1 - t0), "move_evals_per_s": evals / (t3 - t2),
                      "best_key": int(k) ^ (1 << 63)}))   # bench_10k.py prints keys in the signed-order form it exchanges


if __name__ == "__main__":
    main()
