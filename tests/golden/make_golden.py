"""Freezes golden vectors by running the REFERENCE's own code (imported by file path, network stubbed).

Run in the build container only (needs /root/reference):   python tests/golden/make_golden.py
Outputs (committed): tests/golden/greedy_cases.json, p2p_cases.json, eta_feature_cases.json.

What is pinned:
  greedy  — utils._multi_stop (utils.py:85-193): trips, optimized_order, trip count, on integer / float /
            asymmetric / tied matrices, with and without vehicle_capacity / maximum_distance / payload.
  p2p     — utils._point_to_point (utils.py:53-82): the error strings of the single-stop checks.
  eta     — ml.predict_eta_minutes (ml.py:23-58): the 12-feature row handed to model.predict, its column order
            and dtypes, and the ISO completion timestamp for a model that returns a known constant.
Large cases do not store the matrix; they store how it was made (seed → coordinates → oracle matrix, SPEC §2) and
its sha256, so a drift in the matrix definition is detected rather than silently re-pinned.
"""
from __future__ import annotations

import datetime as dt
import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle as O  # noqa: E402
from oracle import ref_loader as RL  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
BOX = (14.4043, 14.6656, 120.9634, 121.0891)   # frontend/map-app/components/MapView.jsx:25
DEPOT = (14.5836, 121.0409)                    # LocationsTableSeeder.php:14


def synth_instance(seed: int, n_stops: int, vehicles: int):
    """SURVEY.md §8(d) synthetic inputs."""
    rng = np.random.default_rng(seed)
    lat = np.concatenate([[DEPOT[0]], rng.uniform(BOX[0], BOX[1], n_stops)])
    lon = np.concatenate([[DEPOT[1]], rng.uniform(BOX[2], BOX[3], n_stops)])
    demand = np.concatenate([[0], rng.integers(1, 10, n_stops)]).astype(np.int64)
    cap = int(np.ceil(1.05 * demand.sum() / vehicles))
    return lat, lon, demand, cap


def run_reference(matrix, payloads, driver_details):
    n = len(matrix)
    source = {"lat": 0.0, "lon": 0.0}
    dests = []
    for i in range(1, n):
        d = {"lat": float(i), "lon": float(i)}
        if payloads[i] is not None:
            d["payload"] = payloads[i]
        dests.append(d)
    feat, stub = RL.reference_trips(matrix, source, dests, driver_details)
    trips = []
    for url, body in stub.calls:
        if "/directions/" in url:
            trips.append([int(c[0]) for c in body["coordinates"]][1:-1])
    props = feat["properties"]
    return {"trips": trips, "optimized_order": props["optimized_order"], "n_trips": props["summary"]["trips"]}


def greedy_cases():
    cases = []

    def add(name, matrix, payloads, dd, note, recipe=None):
        res = run_reference(matrix, payloads, dd)
        c = {"name": name, "note": note, "driver_details": dd, "payloads": payloads, **res}
        if recipe is None:
            c["matrix"] = matrix
        else:
            arr = np.asarray(matrix, dtype=np.int32)
            c["recipe"] = recipe
            c["matrix_sha256"] = hashlib.sha256(arr.tobytes()).hexdigest()
        cases.append(c)

    # the reference's only fixture: tests/test_get_route.py:3-12 (cap 5, two payload-4 stops => 2 trips)
    add("ref_fixture_two_trips", [[0, 4500, 3100], [4500, 0, 1900], [3100, 1900, 0]], [None, 4, 4],
        {"driver_name": "John Doe", "vehicle_type": "car", "vehicle_capacity": 5, "maximum_distance": 50000},
        "request body of tests/test_get_route.py; matrix values are a stand-in")
    # ties resolve to ascending index (stable sort over an ascending list)
    add("ties_ascending_index", [[0, 7, 7, 7, 5], [7, 0, 1, 1, 1], [7, 1, 0, 1, 1], [7, 1, 1, 0, 1], [5, 1, 1, 1, 0]],
        [None, 1, 1, 1, 1], {"vehicle_capacity": 10}, "equal depot distances")
    # a stop skipped for capacity does not end the trip (a stop that can never fit would spin forever in the
    # reference, utils.py:118-139, so every case here is feasible stop by stop)
    add("capacity_skip_continues", [[0, 1, 2, 3, 4], [1, 0, 1, 2, 3], [2, 1, 0, 1, 2], [3, 2, 1, 0, 1], [4, 3, 2, 1, 0]],
        [None, 3, 5, 2, 1], {"vehicle_capacity": 6}, "stop 2 skipped in trip 1, stops 3 and 4 still accepted")
    # maximum_distance binds; explicit 0 capacity is not 'absent' is untestable (would hang) so only max_dist here
    add("max_distance_binds", [[0, 10, 20, 25], [10, 0, 10, 20], [20, 10, 0, 10], [25, 20, 10, 0]],
        [None, 1, 1, 1], {"maximum_distance": 52}, "trip closes when going on would exceed the return budget")
    # no driver details at all: defaults 9e12 / 9e12, payload absent => 0
    add("all_defaults_one_trip", [[0, 3, 1, 2], [3, 0, 5, 5], [1, 5, 0, 5], [2, 5, 5, 0]], [None, None, None, None], {},
        "utils.py:112-113 defaults, utils.py:126 payload default")
    # asymmetric matrix: D[cur][idx] and D[idx][0] are read in that orientation (utils.py:128)
    add("asymmetric", [[0, 5, 6, 7], [9, 0, 1, 9], [1, 9, 0, 1], [2, 1, 9, 0]], [None, 1, 1, 1],
        {"vehicle_capacity": 2, "maximum_distance": 20}, "directional matrix")
    # float matrix and float payloads (ORS returns floats; FP64 accumulation order matters)
    rng = np.random.default_rng(42)
    n = 13
    M = np.round(rng.uniform(100, 9000, (n, n)), 1)
    np.fill_diagonal(M, 0.0)
    add("float_matrix_float_payload", M.tolist(), [None] + [float(x) for x in np.round(rng.uniform(0.1, 2.5, n - 1), 2)],
        {"vehicle_capacity": 4.2, "maximum_distance": 30000.5}, "FP64 path, non-integer everything")

    # BASELINE config 1: 100 stops / 10 vehicles (matrix from SPEC §2, decimetres)
    for name, seed, stops, veh in (("config1_100stops_10veh", 1234, 100, 10), ("config2_1000stops_10veh", 2345, 1000, 10)):
        lat, lon, demand, cap = synth_instance(seed, stops, veh)
        xyz = O.unit_vectors(lat, lon)
        D = O.matrix_rows(xyz, 0, stops + 1, 10.0, 1.0, want=("dist",))["dist"]
        add(name, D.tolist(), [None] + [int(x) for x in demand[1:]], {"vehicle_capacity": cap},
            "SURVEY §8(d) synthetic instance", recipe={"seed": seed, "stops": stops, "vehicles": veh, "dist_scale": 10.0})
    # same instance with a binding maximum_distance (decimetres)
    lat, lon, demand, cap = synth_instance(1234, 100, 10)
    xyz = O.unit_vectors(lat, lon)
    D = O.matrix_rows(xyz, 0, 101, 10.0, 1.0, want=("dist",))["dist"]
    add("config1_maxdist_600km", D.tolist(), [None] + [int(x) for x in demand[1:]],
        {"vehicle_capacity": cap, "maximum_distance": 600000}, "capacity and maximum_distance both bind",
        recipe={"seed": 1234, "stops": 100, "vehicles": 10, "dist_scale": 10.0})
    return cases


def p2p_cases():
    out = []
    for payload, dd, dist in ((4, {"vehicle_capacity": 5, "maximum_distance": 50000}, 6983.0),
                              (6, {"vehicle_capacity": 5, "maximum_distance": 50000}, 6983.0),
                              (4, {"vehicle_capacity": 5, "maximum_distance": 5000}, 6983.0),
                              (6, {"vehicle_capacity": 5, "maximum_distance": 5000}, 6983.0),
                              (10 ** 6, {}, 1e13), (999999, {}, 9e12), (None, {}, 12.5)):
        mod, stub = RL.load_utils(None)

        def post(url, json=None, headers=None, timeout=None, _d=dist):
            feat = {"type": "Feature", "geometry": {"type": "LineString", "coordinates": json["coordinates"]},
                    "properties": {"segments": [], "summary": {"distance": _d, "duration": 1.0}}}
            return RL._Resp({"features": [feat]})

        mod.requests.post = staticmethod(post)
        dest = {"lat": 1.0, "lon": 1.0}
        if payload is not None:
            dest["payload"] = payload
        res = mod._point_to_point({"lat": 0.0, "lon": 0.0}, dest, "driving-car", dd)
        out.append({"payload": payload, "driver_details": dd, "distance": dist, "error": res.get("error")})
    return out


class _FakeModel:
    def __init__(self, value):
        self.value = value
        self.rows = []

    def predict(self, df):
        self.rows.append({"columns": list(df.columns), "dtypes": [str(t) for t in df.dtypes],
                          "values": [float(v) for v in df.iloc[0].tolist()]})
        return [self.value]


def eta_cases():
    ml = RL.load_ml()
    out = []
    inputs = [
        dict(weather="Sunny", traffic="Low", distance_m=25791.9, pickup_time="2025-08-12T14:30:00", driver_age=30.0),
        dict(weather="Stormy", traffic="Jam", distance_m=6983.0, pickup_time="2025-08-17T23:59:59", driver_age=41),
        dict(weather="Cloudy", traffic="High", distance_m=0, pickup_time="2024-02-29T00:00:00", driver_age=0),
        dict(weather="Windy", traffic="Medium", distance_m=None, pickup_time="2025-01-06T08:15:00", driver_age=None),
        dict(weather="Fog", traffic="Unknown", distance_m=123456.789, pickup_time="2025-12-31T12:00:00", driver_age=22.5),
        dict(weather="sunny", traffic="low", distance_m=1.0, pickup_time="2025-03-05T05:05:05", driver_age=19.0),
    ]
    for kw in inputs:
        fm = _FakeModel(37.25)
        ml._model = fm
        minutes, ts = ml.predict_eta_minutes(**kw)
        out.append({"inputs": kw, "row": fm.rows[0], "eta_minutes": minutes, "eta_ts": ts})
    # datetime passthrough
    fm = _FakeModel(12.5)
    ml._model = fm
    minutes, ts = ml.predict_eta_minutes(weather="Sunny", traffic="Low", distance_m=1500.0,
                                         pickup_time=dt.datetime(2025, 8, 12, 9, 0, 0))
    out.append({"inputs": {"weather": "Sunny", "traffic": "Low", "distance_m": 1500.0,
                           "pickup_time": "datetime(2025,8,12,9,0,0)"}, "row": fm.rows[0],
                "eta_minutes": minutes, "eta_ts": ts})
    # model without .predict => (None, None)  (ml.py:25-26)
    ml._model = "ERROR:missing"
    out.append({"inputs": "model unavailable", "result": list(ml.predict_eta_minutes(
        weather="Sunny", traffic="Low", distance_m=1.0, pickup_time=None))})
    return out


def optimize_route_cases():
    """Whole-function goldens: utils.optimize_route (utils.py:10-48) on request bodies, ORS stubbed with a fixed
    matrix and oracle.ref_loader.stub_directions; the stored value is the reference's complete return dict."""
    out = []

    def run(name, body, matrix):
        mod, stub = RL.load_utils(matrix)
        res = mod.optimize_route(json.loads(json.dumps(body)))
        out.append({"name": name, "body": body, "matrix": matrix, "result": res})

    fixture = {"source_point": {"lat": 14.584630, "lon": 121.056885},                     # tests/test_get_route.py:3-12
               "destination_points": [{"lat": 14.544145, "lon": 121.056617, "payload": 4},
                                      {"lat": 14.557855, "lon": 121.066139, "payload": 4}],
               "driver_details": {"driver_name": "John Doe", "vehicle_type": "car", "vehicle_capacity": 5,
                                  "maximum_distance": 50000}}
    run("ref_fixture", fixture, [[0, 4500.5, 3100.2], [4510.0, 0, 1900.7], [3050.9, 1950.1, 0]])
    run("empty_body", {}, None)
    run("no_destinations", {"source_point": {"lat": 1, "lon": 2}, "destination_points": []}, None)
    one = {"source_point": {"lat": 14.58, "lon": 121.05}, "destination_points": [{"lat": 14.55, "lon": 121.06, "payload": 3}],
           "driver_details": {"vehicle_type": " Truck ", "vehicle_capacity": 5, "driver_name": "A"}}
    run("single_stop_ok", one, None)
    one_bad = json.loads(json.dumps(one))
    one_bad["driver_details"].update(vehicle_capacity=2, maximum_distance=100)
    run("single_stop_both_errors", one_bad, None)
    rng = np.random.default_rng(5)
    n = 9
    M = np.round(rng.uniform(500, 9000, (n, n)), 1)
    np.fill_diagonal(M, 0.0)
    body = {"source_point": {"lat": 14.58, "lon": 121.04},
            "destination_points": [{"lat": round(14.45 + 0.02 * i, 4), "lon": round(120.98 + 0.01 * i, 4),
                                    "payload": int(rng.integers(1, 5))} for i in range(n - 1)],
            "driver_details": {"vehicle_type": "motorcycle", "vehicle_capacity": 7, "maximum_distance": 26000.0},
            "meta": {"origin_id": 1}, "use_ml_eta": False}
    run("eight_stops_float_matrix", body, M.tolist())
    body2 = json.loads(json.dumps(body))
    body2["driver_details"] = {}
    for d in body2["destination_points"][::2]:
        d.pop("payload")
    run("eight_stops_defaults_missing_payloads", body2, M.tolist())
    run("matrix_missing", body, [])
    body3 = json.loads(json.dumps(body))                      # unknown vehicle type -> default profile (utils.py:22-29)
    body3["driver_details"] = {"vehicle_type": "Hovercraft", "vehicle_capacity": 9, "driver_name": None}
    run("eight_stops_unknown_vehicle", body3, M.tolist())
    body4 = json.loads(json.dumps(body))                      # nothing to carry, nothing to carry it in: 0 <= 0 holds
    for d in body4["destination_points"]:
        d["payload"] = 0
    body4["driver_details"] = {"vehicle_type": "bike", "vehicle_capacity": 0}
    run("eight_stops_zero_payload_zero_capacity", body4, M.tolist())
    Mi = np.round(M)                                          # limits met with equality: both tests are "<=" (utils.py:127-129)
    body5 = json.loads(json.dumps(body))
    far = float(max(Mi[0, i] + Mi[i, 0] for i in range(1, n)))   # the longest there-and-back alone exactly fits
    body5["driver_details"] = {"vehicle_type": "car", "maximum_distance": far,
                               "vehicle_capacity": max(d["payload"] for d in body5["destination_points"])}
    run("eight_stops_limits_met_with_equality", body5, Mi.tolist())
    return out


def main():
    assert RL.available(), "run this where /root/reference exists"
    for name, fn in (("greedy_cases.json", greedy_cases), ("p2p_cases.json", p2p_cases),
                     ("eta_feature_cases.json", eta_cases), ("optimize_route_cases.json", optimize_route_cases)):
        with open(os.path.join(HERE, name), "w") as f:
            json.dump(fn(), f, indent=None, separators=(",", ":"))
            f.write("\n")
        print("wrote", name, os.path.getsize(os.path.join(HERE, name)), "bytes")


if __name__ == "__main__":
    main()
