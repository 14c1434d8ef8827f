import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")

BOX = (14.4043, 14.6656, 120.9634, 121.0891)  # Metro Manila, SURVEY.md §8(d)
DEPOT = (14.5836, 121.0409)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


def synth_instance(seed: int, n_stops: int, vehicles: int):
    """SURVEY.md §8(d) synthetic instance (same recipe as tests/golden/make_golden.py)."""
    rng = np.random.default_rng(seed)
    lat = np.concatenate([[DEPOT[0]], rng.uniform(BOX[0], BOX[1], n_stops)])
    lon = np.concatenate([[DEPOT[1]], rng.uniform(BOX[2], BOX[3], n_stops)])
    demand = np.concatenate([[0], rng.integers(1, 10, n_stops)]).astype(np.int64)
    cap = int(np.ceil(1.05 * demand.sum() / vehicles))
    return lat, lon, demand, cap


def load_golden(name):
    with open(os.path.join(GOLDEN, name)) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def engine():
    from routest_b200 import Engine
    eng = Engine(0)  # raises if the library or the GPU is missing: GPU tests must not pass on a fallback
    yield eng
    eng.close()

TS_CAR = 3.6 / 40.0  # driving-car seconds per metre (SPEC §3)


def synth_time_windows(seed: int, t_out, t_back, half_lo=900, half_hi=3600, horizon=43200, service=300):
    """centre uniform in the horizon, half-width U{half_lo..half_hi} s, clipped; depot [0, horizon].
    t_out/t_back are the depot->stop / stop->depot travel times: the centre is clamped into the span in which the
    stop can be served at all, so that every stop is feasible on its own (the reference would hang otherwise)."""
    n_stops = len(t_out)
    rng = np.random.default_rng(seed + 7919)
    centre = rng.integers(0, horizon + 1, n_stops)
    half = rng.integers(half_lo, half_hi + 1, n_stops)
    centre = np.clip(centre, t_out, horizon - service - np.asarray(t_back))
    ready = np.clip(centre - half, 0, horizon)
    due = np.clip(centre + half, 0, horizon)
    tw = np.zeros((n_stops + 1, 3), np.int32)
    tw[0] = (0, horizon, 0)
    tw[1:, 0] = ready
    tw[1:, 1] = due
    tw[1:, 2] = service
    return tw


def synth_vrptw(seed: int, n_stops: int, vehicles: int, **tw_kw):
    """A full VRPTW instance on the oracle's matrix plus a feasible start from the TW-aware greedy (SPEC §4).
    Returns (oracle Instance, node[] positions, dict of raw arrays)."""
    from oracle import oracle as O
    lat, lon, demand, cap = synth_instance(seed, n_stops, vehicles)
    n = n_stops + 1
    m = O.matrix_rows(O.unit_vectors(lat, lon), 0, n, 10.0, TS_CAR, threads=O.num_threads())
    tw = synth_time_windows(seed, m["time"][0, 1:], m["time"][1:, 0], **tw_kw)
    stops = np.zeros((n, 4), np.int32)
    stops[:, :3] = tw
    stops[:, 3] = demand
    rc, trips = O.greedy_trips(m["dist"], demand.astype(np.float64), float(cap), 9e12, T=m["time"], tw=tw)
    assert rc == O.OK, "synthetic instance has a stop that is infeasible on its own"
    inst = O.Instance(m["dist"], m["time"], stops, cap)
    node = O.build_positions(trips, n)
    return inst, node, dict(lat=lat, lon=lon, demand=demand, cap=cap, tw=tw, trips=trips)
