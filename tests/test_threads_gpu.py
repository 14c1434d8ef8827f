"""GPU: the host-buffer entry points are safe to call from several threads on one engine (the reference's handlers
run on arbitrary Flask worker threads, app.py:8 `threaded=True`; SURVEY.md §8b asks for a re-entrant C ABI).  ctypes
drops the GIL during a call, so the threads really meet inside the library; calls on one engine are serialised
there and every thread must get the result it would get alone."""
import threading

import numpy as np
import pytest

from conftest import TS_CAR, synth_instance, synth_vrptw
from oracle import oracle as O

pytestmark = pytest.mark.gpu


def test_concurrent_callers_get_their_own_results(engine):
    cases = [synth_instance(700 + k, 60 + 25 * k, 4 + k) for k in range(6)]
    want = []
    for lat, lon, demand, cap in cases:
        m = O.matrix_rows(O.unit_vectors(lat, lon), 0, lat.size, 10.0, TS_CAR)
        rc, trips = O.greedy_trips(m["dist"], demand.astype(float), float(cap), 9e12)
        assert rc == 0
        want.append((m["dist"], m["time"], trips))
    errors, rounds = [], 6

    def worker(k):
        try:
            lat, lon, demand, cap = cases[k]
            for _ in range(rounds):
                got = engine.matrix_build(lat, lon, dist_scale=10.0, time_scale=TS_CAR)
                assert np.array_equal(got["dist"], want[k][0]) and np.array_equal(got["time"], want[k][1])
                trips, infeasible = engine.greedy_trips(got["dist"].astype(np.float64), demand.astype(float), float(cap),
                                                        9e12)
                assert not infeasible and trips == want[k][2]
        except Exception as exc:  # noqa: BLE001 - reported by the main thread
            errors.append((k, repr(exc)))

    threads = [threading.Thread(target=worker, args=(k,)) for k in range(len(cases))]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors


def test_problems_of_one_engine_from_different_threads(engine):
    insts = [synth_vrptw(800 + k, 70 + 20 * k, 5) for k in range(4)]
    refs = [O.local_search(inst, node) for inst, node, _ in insts]
    errors = []

    def worker(k):
        try:
            inst, node, raw = insts[k]
            with engine.problem(inst.n, 1) as pr:
                pr.set_coords(raw["lat"][None], raw["lon"][None])
                pr.build_matrix(10.0, TS_CAR)
                pr.set_stops(inst.stops[None], [inst.cap], [inst.max_dist])
                pr.set_routes([O.positions_to_trips(node)])
                moves, cost, keys = pr.local_search(keys_capacity=4 * inst.n)
                assert np.array_equal(keys[0], refs[k][1]) and cost[0] == refs[k][2]
                assert np.array_equal(pr.get_state(0)["node"], refs[k][0])
        except Exception as exc:  # noqa: BLE001
            errors.append((k, repr(exc)))

    threads = [threading.Thread(target=worker, args=(k,)) for k in range(len(insts))]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors
