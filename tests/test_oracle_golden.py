"""CPU: the oracle against the golden vectors frozen from the reference's own code, and against libm."""
import hashlib

import numpy as np
import pytest

from conftest import load_golden, synth_instance
from oracle import oracle as O


def _case_matrix(case):
    if "matrix" in case:
        return np.array(case["matrix"], dtype=np.float64)
    r = case["recipe"]
    lat, lon, _, _ = synth_instance(r["seed"], r["stops"], r["vehicles"])
    D = O.matrix_rows(O.unit_vectors(lat, lon), 0, r["stops"] + 1, r["dist_scale"], 1.0, want=("dist",))["dist"]
    assert hashlib.sha256(D.tobytes()).hexdigest() == case["matrix_sha256"], "SPEC §2 matrix drifted from the golden"
    return D


@pytest.mark.parametrize("case", load_golden("greedy_cases.json"), ids=lambda c: c["name"])
def test_greedy_matches_reference(case):
    """or_greedy_trips == reference _multi_stop (utils.py:111-139) on the frozen cases."""
    D = _case_matrix(case)
    dd = case["driver_details"]
    demand = [0.0] + [float(p) if p is not None else 0.0 for p in case["payloads"][1:]]
    cap = float(dd.get("vehicle_capacity", 9e12))      # utils.py:112
    max_dist = float(dd.get("maximum_distance", 9e12))  # utils.py:113
    rc, trips = O.greedy_trips(D, demand, cap, max_dist)
    assert rc == O.OK
    assert trips == case["trips"]
    assert len(trips) == case["n_trips"]
    assert [i - 1 for t in trips for i in t] == case["optimized_order"]  # utils.py:172-175


def test_greedy_infeasible_stop_terminates():
    """The reference spins forever on a stop that fits no trip; SPEC §4 defines an error instead."""
    D = [[0, 1, 2], [1, 0, 1], [2, 1, 0]]
    rc, trips = O.greedy_trips(D, [0, 3, 9], 6, 9e12)
    assert rc == O.ERR_INFEASIBLE_STOP
    assert trips == [[1]]


@pytest.mark.parametrize("case", load_golden("p2p_cases.json"))
def test_p2p_check_matches_reference(case):
    dd = case["driver_details"]
    payload = case["payload"] if case["payload"] is not None else 0   # utils.py:69
    cap = dd.get("vehicle_capacity", 999999)                          # utils.py:70
    max_dist = float(dd.get("maximum_distance", 9e12))                # utils.py:71
    mask = O.p2p_check(payload, cap, case["distance"], max_dist)
    msgs = []
    if mask & 1:
        msgs.append("payload exceeds vehicle capacity")
    if mask & 2:
        msgs.append("route distance exceeds maximum_distance")
    assert (" | ".join(msgs) or None) == case["error"]


def test_sincos_accuracy():
    t = np.linspace(-3.2, 3.2, 400001)
    s, c = O.sincos(t)
    assert np.abs(s - np.sin(t)).max() <= 3e-16
    assert np.abs(c - np.cos(t)).max() <= 3e-16


def test_matrix_against_libm_haversine_metro():
    lat, lon, _, _ = synth_instance(1234, 100, 10)
    xyz = O.unit_vectors(lat, lon)
    r = O.matrix_rows(xyz, 0, 101, 10.0, 0.09, want=("dist", "time", "metres"))
    h = O.haversine_libm(lat, lon)
    off = h > 0
    # FP64 output: north-star bound is 1e-6 relative; SPEC promises far better
    assert (np.abs(r["metres"] - h)[off] / h[off]).max() < 1e-9
    # integer cells: rounding of a value whose relative error is <= 3e-7
    assert (np.abs(r["dist"] - h * 10.0) <= 0.5 + 1e-6 * h * 10.0).all()
    assert (np.abs(r["time"] - h * 0.09) <= 0.5 + 1e-6 * h * 0.09).all()
    assert (r["dist"] == r["dist"].T).all() and (np.diag(r["dist"]) == 0).all()


def test_matrix_against_libm_haversine_global():
    rng = np.random.default_rng(7)
    lat = np.concatenate([rng.uniform(-90, 90, 200), [90.0, -90.0, 0.0, 0.0, 0.0]])
    lon = np.concatenate([rng.uniform(-180, 180, 200), [0.0, 0.0, 180.0, -180.0, 0.0]])
    xyz = O.unit_vectors(lat, lon)
    r = O.matrix_rows(xyz, 0, lat.size, 0.01, 0.001, want=("dist", "metres"))
    h = O.haversine_libm(lat, lon)
    off = h > 1.0
    assert (np.abs(r["metres"] - h)[off] / h[off]).max() < 1e-9
    assert (np.abs(r["dist"] - h * 0.01) <= 0.5 + 1e-6 * h * 0.01).all()


def test_matrix_row_range():
    lat, lon, _, _ = synth_instance(99, 40, 4)
    xyz = O.unit_vectors(lat, lon)
    full = O.matrix_rows(xyz, 0, 41, 10.0, 0.09)
    part = O.matrix_rows(xyz, 7, 19, 10.0, 0.09)
    assert (part["dist"] == full["dist"][7:19]).all() and (part["time"] == full["time"][7:19]).all()
