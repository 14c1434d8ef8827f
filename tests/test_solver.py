"""The solver front end (routest_b200/solver.py) against whole-function goldens of the reference's
utils.optimize_route (utils.py:10-48) and, for the network-free default path, against the oracle."""
import json

import numpy as np
import pytest

from conftest import TS_CAR, load_golden, synth_instance
from oracle import oracle as O
from oracle.ref_loader import stub_directions
from routest_b200 import solver as S

CASES = {c["name"]: c for c in load_golden("optimize_route_cases.json")}


def _hooks(case):
    matrix = case["matrix"]
    return dict(matrix_fn=lambda coords, profile: {"distances": matrix},
                directions_fn=lambda coords, profile: stub_directions(coords), engine_label="backend:ors")


@pytest.mark.parametrize("name", ["empty_body", "no_destinations", "single_stop_ok", "single_stop_both_errors"])
def test_host_only_paths_equal_reference(name):
    """No matrix, no trips: validation (utils.py:15-16) and the single-stop checks (utils.py:68-80)."""
    case = CASES[name]
    assert S.optimize_route(json.loads(json.dumps(case["body"])), **_hooks(case)) == case["result"]


def test_missing_source_point_raises_like_the_reference():
    with pytest.raises(KeyError):  # utils.py:31 indexes input_data["source_point"]
        S.optimize_route({"destination_points": [{"lat": 1, "lon": 2}]}, **_hooks(CASES["single_stop_ok"]))


def test_matrix_service_without_distances():
    case = CASES["matrix_missing"]
    assert S.optimize_route(json.loads(json.dumps(case["body"])), **_hooks(case)) == case["result"]


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["ref_fixture", "eight_stops_float_matrix", "eight_stops_defaults_missing_payloads",
                                  "eight_stops_unknown_vehicle", "eight_stops_zero_payload_zero_capacity",
                                  "eight_stops_limits_met_with_equality"])
def test_multi_stop_equals_reference_whole_feature(engine, name):
    """Same request, same matrix answer, same directions answers => the reference's complete return dict."""
    case = CASES[name]
    got = S.optimize_route(json.loads(json.dumps(case["body"])), engine=engine, **_hooks(case))
    assert got == case["result"]


@pytest.mark.gpu
def test_infeasible_stop_is_reported_not_hung(engine):
    body = json.loads(json.dumps(CASES["eight_stops_float_matrix"]["body"]))
    body["driver_details"]["vehicle_capacity"] = 1
    res = S.optimize_route(body, engine=engine, **_hooks(CASES["eight_stops_float_matrix"]))
    assert "error" in res and "fit no trip" in res["error"]


def _request(seed, stops, veh, **extra):
    lat, lon, demand, cap = synth_instance(seed, stops, veh)
    body = {"source_point": {"lat": float(lat[0]), "lon": float(lon[0])},
            "destination_points": [{"lat": float(lat[i]), "lon": float(lon[i]), "payload": int(demand[i])}
                                   for i in range(1, stops + 1)],
            "driver_details": {"driver_name": "D", "vehicle_type": "car", "vehicle_capacity": cap, **extra}}
    return body, lat, lon, demand, cap


@pytest.mark.gpu
def test_offline_default_path_shape_and_trips(engine):
    """No hooks: engine matrix + engine greedy + straight-line directions.  Trips equal the oracle greedy on the
    oracle matrix; the Feature has every field the frontend reads."""
    body, lat, lon, demand, cap = _request(1234, 100, 10)
    res = S.optimize_route(body, engine=engine)
    m = O.matrix_rows(O.unit_vectors(lat, lon), 0, 101, 10.0, TS_CAR)
    rc, trips = O.greedy_trips(m["dist"], demand.astype(float), float(cap), 9e12 * 10)
    assert rc == 0
    p = res["properties"]
    assert p["optimized_order"] == [i - 1 for t in trips for i in t]
    assert p["summary"]["trips"] == len(trips) and p["engine"] == "backend:b200" and p["vehicle_type"] == "car"
    want_d = sum(m["dist"][a, b] for t in trips for a, b in zip([0] + t, t + [0])) / 10.0
    want_t = sum(int(m["time"][a, b]) for t in trips for a, b in zip([0] + t, t + [0]))
    assert abs(p["summary"]["distance"] - want_d) < 1e-6 and p["summary"]["duration"] == want_t
    assert res["type"] == "Feature" and res["geometry"]["type"] == "LineString"
    assert len(res["geometry"]["coordinates"]) == 100 + 2 * len(trips)
    assert len(res["bbox"]) == 4 and p["destinations"] == body["destination_points"] and p["source"] == body["source_point"]
    step = p["segments"][0]["steps"][0]
    assert {"instruction", "type", "name", "distance", "duration"} <= set(step)
    json.dumps(res)  # plain JSON types only


@pytest.mark.gpu
def test_single_stop_offline(engine):
    body, *_ = _request(77, 1, 1)
    res = S.optimize_route(body, engine=engine)
    assert res["properties"]["optimized_order"] == [0] and res["properties"]["summary"]["distance"] > 0
    body["driver_details"]["maximum_distance"] = 1.0
    assert S.optimize_route(body, engine=engine) == {"error": "route distance exceeds maximum_distance"}


@pytest.mark.gpu
def test_time_windows_and_improve_extension(engine):
    """Optional fields: windows respected by the construction, local search never worsens the distance."""
    body, lat, lon, demand, cap = _request(4321, 120, 8)
    m = O.matrix_rows(O.unit_vectors(lat, lon), 0, 121, 10.0, TS_CAR)
    rng = np.random.default_rng(1)
    for i, d in enumerate(body["destination_points"], start=1):
        c = int(np.clip(rng.integers(0, 43200), m["time"][0, i], 43200 - 300 - m["time"][i, 0]))
        d.update(tw_start=max(0, c - 3000), tw_end=min(43200, c + 3000), service_time=300)
    base = S.optimize_route(json.loads(json.dumps(body)), engine=engine)
    body["driver_details"]["improve"] = True
    better = S.optimize_route(json.loads(json.dumps(body)), engine=engine)
    assert "error" not in base and "error" not in better
    assert sorted(base["properties"]["optimized_order"]) == list(range(120))
    assert sorted(better["properties"]["optimized_order"]) == list(range(120))
    assert better["properties"]["summary"]["distance"] < base["properties"]["summary"]["distance"]
    body["driver_details"]["improve"] = "batched"                      # SPEC §8.1: sets of independent moves
    batched = S.optimize_route(json.loads(json.dumps(body)), engine=engine)
    assert sorted(batched["properties"]["optimized_order"]) == list(range(120))
    assert batched["properties"]["summary"]["distance"] < base["properties"]["summary"]["distance"]
    # the greedy with windows equals the oracle's
    stops = np.zeros((121, 3), np.int32)
    stops[0] = (0, 43200, 0)
    for i, d in enumerate(body["destination_points"], start=1):
        stops[i] = (d["tw_start"], d["tw_end"], d["service_time"])
    rc, trips = O.greedy_trips(m["dist"], demand.astype(float), float(cap), 9e13, T=m["time"], tw=stops)
    assert rc == 0 and base["properties"]["optimized_order"] == [i - 1 for t in trips for i in t]


@pytest.mark.gpu
def test_improve_on_a_directional_service_matrix(engine):
    """Optional fields with a matrix hook: the service's (asymmetric) matrix is integerised, windows are honoured
    and the local search runs on it; same trips as the oracle on the same integer matrix."""
    rng = np.random.default_rng(3)
    n = 41
    base = rng.uniform(500, 12000, (n, n))
    D = np.round(base + rng.uniform(0, 900, (n, n)), 1)        # directional
    np.fill_diagonal(D, 0.0)
    body = {"source_point": {"lat": 14.58, "lon": 121.04},
            "destination_points": [{"lat": 14.40 + 0.005 * i, "lon": 120.97 + 0.003 * i, "payload": int(rng.integers(1, 5)),
                                    "tw_start": 0, "tw_end": 43200, "service_time": 120} for i in range(n - 1)],
            "driver_details": {"vehicle_type": "car", "vehicle_capacity": 18, "improve": True}}
    hooks = dict(matrix_fn=lambda coords, profile: {"distances": D.tolist()},
                 directions_fn=lambda coords, profile: stub_directions(coords))
    res = S.optimize_route(json.loads(json.dumps(body)), engine=engine, **hooks)
    assert "error" not in res and sorted(res["properties"]["optimized_order"]) == list(range(n - 1))
    Di = np.rint(D * 10).astype(np.int32)
    Ti = np.rint(D * TS_CAR).astype(np.int32)
    demand = np.array([0] + [d["payload"] for d in body["destination_points"]])
    stops = np.zeros((n, 4), np.int32)
    stops[:, 1] = 43200
    stops[1:, 2] = 120
    stops[:, 3] = demand
    rc, trips = O.greedy_trips(Di, demand.astype(float), 18.0, 9e13, T=Ti, tw=stops[:, :3])
    inst = O.Instance(Di, Ti, stops, 18)
    final, _, _ = O.local_search(inst, O.build_positions(trips, n))
    want = [i - 1 for t in O.positions_to_trips(final) if t for i in t]
    assert rc == 0 and res["properties"]["optimized_order"] == want


@pytest.mark.gpu
def test_engine_failures_come_back_as_error_dicts(engine):
    """utils.py:63-66,106-109: a failed matrix/directions call is returned as {"error": ...}, never raised; the same
    holds for a failed engine call (here: a latitude the engine refuses)."""
    body, *_ = _request(99, 12, 3)
    body["destination_points"][4]["lat"] = 123.0
    res = S.optimize_route(body, engine=engine)
    assert set(res) == {"error"} and "route engine failed" in res["error"] and "coordinate" in res["error"]
