"""Times the REFERENCE's own greedy (Flaskr/utils.py:_multi_stop, imported by file path, ORS stubbed) on synthetic
instances — the "reference code" column of SURVEY.md §8(d).  Runs only where /root/reference exists (the build
container, no GPU needed): single Python thread, as the reference is per request.  The matrix is the CPU oracle's
great-circle matrix in metres handed over as list-of-lists of floats, exactly what the stubbed ORS call returns.

  python tests/perf/reference_greedy_time.py
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from bench import synth_coords  # noqa: E402
from oracle import oracle as O, ref_loader  # noqa: E402


def main():
    if not ref_loader.available():
        print(json.dumps({"unavailable": "/root/reference is not present on this machine"}))
        return
    for stops, veh in ((100, 10), (1000, 10), (5000, 50)):
        n = stops + 1
        lat, lon, demand, cap = synth_coords(2345, stops, veh)
        m = O.matrix_rows(O.unit_vectors(lat, lon), 0, n, 1.0, 0.09, threads=O.num_threads())
        D = m["dist"].astype(float).tolist()
        src = {"lat": float(lat[0]), "lon": float(lon[0])}
        dests = [{"lat": float(lat[i]), "lon": float(lon[i]), "payload": int(demand[i])} for i in range(1, n)]
        details = {"vehicle_capacity": int(cap)}
        mod, stub = ref_loader.load_utils(D)                    # module import and stub set-up stay outside the clock
        t0 = time.perf_counter()
        feat = mod._multi_stop(src, dests, "driving-car", details)
        dt = time.perf_counter() - t0
        trips = sum(1 for url, _ in stub.calls if "/directions/" in url)
        # the oracle (C restatement of the same loop) on the same matrix, for scale
        t1 = time.perf_counter()
        rc, otrips = O.greedy_trips(m["dist"], demand.astype(float), float(cap), 9e12)
        dc = time.perf_counter() - t1
        order = feat["properties"]["optimized_order"]
        print(json.dumps({"stops": stops, "trips": trips, "reference_python_ms": round(dt * 1e3, 2),
                          "reference_stops_per_s": round(stops / dt), "oracle_c_ms": round(dc * 1e3, 3),
                          "same_order": order == [i - 1 for t in otrips for i in t]}))


if __name__ == "__main__":
    main()
