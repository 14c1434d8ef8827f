"""Latency of one optimize_route request through the reference-facing Python function (development aid):
BASELINE config 1 (100 stops / 10 vehicles) and a 1,000-stop request, offline directions, one GPU."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import synth_coords  # noqa: E402
from routest_b200 import Engine, solver as S  # noqa: E402


def body(seed, stops, vehicles, **details):
    lat, lon, demand, cap = synth_coords(seed, stops, vehicles)
    return {"source_point": {"lat": float(lat[0]), "lon": float(lon[0])},
            "destination_points": [{"lat": float(lat[i]), "lon": float(lon[i]), "payload": int(demand[i])}
                                   for i in range(1, stops + 1)],
            "driver_details": dict(vehicle_type="car", vehicle_capacity=int(cap), **details)}


def main():
    eng = Engine(0)
    for stops, veh in ((100, 10), (1000, 10)):
        for label, details in (("reference fields only", {}), ("improve", {"improve": True}),
                               ("improve batched", {"improve": "batched"})):
            req = body(1234, stops, veh, **details)
            S.optimize_route(json.loads(json.dumps(req)), engine=eng)          # warm-up
            ts = []
            for _ in range(10):
                r = json.loads(json.dumps(req))
                t0 = time.perf_counter()
                res = S.optimize_route(r, engine=eng)
                ts.append(time.perf_counter() - t0)
            assert "error" not in res
            print(json.dumps({"stops": stops, "request": label, "ms_median": round(1e3 * float(np.median(ts)), 3),
                              "ms_min": round(1e3 * min(ts), 3), "trips": res["properties"]["summary"]["trips"]}))
    eng.close()


if __name__ == "__main__":
    main()
