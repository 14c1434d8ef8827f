"""GPU: BASELINE configs 4 and 5 at full size — 10,000-stop best move and 5,000-stop / 50-vehicle construction plus
the first local-search moves — against the oracle, plus size-independent properties."""
import numpy as np
import pytest

from conftest import TS_CAR, synth_vrptw
from oracle import oracle as O

pytestmark = pytest.mark.gpu


def _problem(engine, inst, raw, node):
    pr = engine.problem(inst.n, 1)
    pr.set_coords(raw["lat"], raw["lon"])
    pr.build_matrix(10.0, TS_CAR)
    pr.set_stops(inst.stops[None], [inst.cap])
    return pr


@pytest.mark.timeout(600)
def test_config5_5000_stops_50_vehicles(engine):
    inst, node, raw = synth_vrptw(5678, 5000, 50)
    with _problem(engine, inst, raw, node) as pr:
        trips, status = pr.greedy(raw["demand"][None].astype(float), [float(raw["cap"])], use_tw=True)
        assert status[0] == 0 and trips[0] == raw["trips"]          # construction == oracle, 5,000 stops
        pr.set_routes(trips)
        st = pr.get_state(0)
        ref = O.route_state(inst, node)
        assert st["feasible"] and st["cost"] == ref["cost"] and np.array_equal(st["B"], ref["B"]) and np.array_equal(st["Z"], ref["Z"])
        moves, cost, keys = pr.local_search(max_moves=40, keys_capacity=64)
        final, okeys, ocost = O.local_search(inst, node, max_moves=40, threads=O.num_threads())
        assert moves[0] == len(okeys) == 40 and np.array_equal(keys[0], okeys) and cost[0] == ocost
        end = pr.get_state(0)
        assert end["feasible"] and np.array_equal(end["node"], final)
        assert sorted(end["node"][end["node"] != 0].tolist()) == list(range(1, inst.n))
        assert cost[0] == ref["cost"] + sum(O.key_unpack(k)[0] for k in okeys)
        # SPEC §8.1 from the same start: three batched iterations, then on to a local optimum
        pr.set_routes(trips)
        bmoves, bcost, bkeys = pr.local_search(max_moves=3, keys_capacity=1024, batched=True)
        bfinal, obkeys, sizes, obcost = O.local_search_batched(inst, node, max_iters=3, threads=O.num_threads())
        assert sizes.size == 3 and sizes.min() > 5
        assert bmoves[0] == obkeys.size and np.array_equal(bkeys[0], obkeys) and bcost[0] == obcost
        assert np.array_equal(pr.get_state(0)["node"], bfinal)
        pr.local_search(batched=True)
        end = pr.get_state(0)
        assert end["feasible"] and sorted(end["node"][end["node"] != 0].tolist()) == list(range(1, inst.n))
        assert O.key_unpack(int(pr.eval_moves()[0]))[0] >= 0        # no improving move is left
        assert end["cost"] < 0.25 * ref["cost"]


@pytest.mark.timeout(600)
def test_config4_10000_stops_best_move_and_shards(engine):
    inst, node, raw = synth_vrptw(4567, 10000, 100)
    with _problem(engine, inst, raw, node) as pr:
        pr.set_routes([raw["trips"]])
        full = int(pr.eval_moves()[0])
        assert full == O.best_move(inst, node, threads=O.num_threads())
        parts = [int(pr.eval_moves(shard=(i, 8))[0]) for i in range(8)]   # what 8 ranks would each find
        assert min(parts) == full
        P = node.size
        d, t = pr.get_position_matrix(0, P)
        idx = np.array([0, 1, 2, 77, P // 2, P - 2, P - 1])
        assert np.array_equal(d[idx], inst.D[node[idx]][:, node]) and np.array_equal(t[idx], inst.T[node[idx]][:, node])
        assert np.array_equal(d, d.T)
        # one batched iteration at this size (262 routes: the multi-kernel selection path)
        bmoves, bcost, bkeys = pr.local_search(max_moves=1, keys_capacity=1024, batched=True)
        bfinal, obkeys, sizes, obcost = O.local_search_batched(inst, node, max_iters=1, threads=O.num_threads())
        assert sizes[0] > 20 and np.array_equal(bkeys[0], obkeys) and bcost[0] == obcost
        assert np.array_equal(pr.get_state(0)["node"], bfinal)
