"""GPU: greedy trip construction through the C ABI — bit-exact against the reference's own _multi_stop
(golden vectors frozen by tests/golden/make_golden.py) and against the oracle."""
import hashlib

import numpy as np
import pytest

from conftest import TS_CAR, load_golden, synth_instance, synth_vrptw
from oracle import oracle as O

pytestmark = pytest.mark.gpu


def _case_matrix(case):
    if "matrix" in case:
        return np.array(case["matrix"], dtype=np.float64)
    r = case["recipe"]
    lat, lon, _, _ = synth_instance(r["seed"], r["stops"], r["vehicles"])
    D = O.matrix_rows(O.unit_vectors(lat, lon), 0, r["stops"] + 1, r["dist_scale"], 1.0, want=("dist",))["dist"]
    assert hashlib.sha256(D.tobytes()).hexdigest() == case["matrix_sha256"]
    return D.astype(np.float64)


@pytest.mark.parametrize("case", load_golden("greedy_cases.json"), ids=lambda c: c["name"])
def test_greedy_matches_reference_golden(engine, case):
    """rt_greedy_trips == reference utils._multi_stop (utils.py:111-139): trips, order and trip count."""
    D = _case_matrix(case)
    dd = case["driver_details"]
    demand = [0.0] + [float(p) if p is not None else 0.0 for p in case["payloads"][1:]]
    trips, infeasible = engine.greedy_trips(D, demand, float(dd.get("vehicle_capacity", 9e12)),
                                            float(dd.get("maximum_distance", 9e12)))
    assert not infeasible
    assert trips == case["trips"]
    assert len(trips) == case["n_trips"]
    assert [i - 1 for t in trips for i in t] == case["optimized_order"]


def test_greedy_infeasible_stop_is_an_error_not_a_hang(engine):
    trips, infeasible = engine.greedy_trips([[0, 1, 2], [1, 0, 1], [2, 1, 0]], [0, 3, 9], 6, 9e12)
    assert infeasible and trips == [[1]]


def test_greedy_single_and_empty(engine):
    assert engine.greedy_trips([[0.0]], [0.0], 1, 1) == ([], False)
    assert engine.greedy_trips([[0, 5], [5, 0]], [0, 1], 1, 10) == ([[1]], False)


@pytest.mark.parametrize("seed,stops,veh", [(1234, 100, 10), (2345, 1000, 10), (31, 257, 7)])
def test_problem_greedy_on_device_matrix(engine, seed, stops, veh):
    """Matrix built on the device, greedy run on it in place: equals the oracle on the oracle's matrix."""
    lat, lon, demand, cap = synth_instance(seed, stops, veh)
    n = stops + 1
    D = O.matrix_rows(O.unit_vectors(lat, lon), 0, n, 10.0, TS_CAR, want=("dist",), threads=O.num_threads())["dist"]
    rc, ref = O.greedy_trips(D, demand.astype(float), float(cap), 9e12)
    with engine.problem(n, 1) as pr:
        pr.set_coords(lat, lon)
        pr.build_matrix(10.0, TS_CAR)
        trips, status = pr.greedy(demand.astype(float), float(cap))
    assert status[0] == 0 and rc == 0
    assert trips[0] == ref


def test_problem_greedy_batch_with_time_windows(engine):
    """Batched instances, time-window extension of SPEC §4 (absent in the reference) against the oracle."""
    B, stops = 5, 120
    n = stops + 1
    insts = [synth_vrptw(500 + b, stops, 6) for b in range(B)]
    with engine.problem(n, B) as pr:
        pr.set_coords(np.stack([r["lat"] for _, _, r in insts]), np.stack([r["lon"] for _, _, r in insts]))
        pr.build_matrix(10.0, TS_CAR)
        pr.set_stops(np.stack([i.stops for i, _, _ in insts]), [r["cap"] for _, _, r in insts])
        trips, status = pr.greedy(np.stack([r["demand"] for _, _, r in insts]).astype(float),
                                  [float(r["cap"]) for _, _, r in insts], use_tw=True)
    for b in range(B):
        assert status[b] == 0
        assert trips[b] == insts[b][2]["trips"]
