"""CPU: the oracle's two evaluations of every candidate — brute force (apply + recompute SPEC §6) and the O(1)/sweep
algebra the engine implements — must agree; best-move keys, candidate counts and local search are consistent."""
import numpy as np
import pytest

from conftest import synth_vrptw
from oracle import oracle as O


def _random_state(inst, node, rng, moves):
    """Walk away from the greedy start through random feasible moves so that routes are not depot-sorted."""
    node = node.copy()
    for _ in range(moves):
        fam = int(rng.integers(0, 3))
        d, f = O.eval_fast(inst, node, fam)
        p, q = np.nonzero(f == 1)
        if p.size == 0:
            continue
        i = int(rng.integers(0, p.size))
        node = O.apply_move(node, fam, int(p[i]), int(q[i]))
        assert O.route_state(inst, node)["feasible"]
    return node


@pytest.mark.parametrize("seed,stops,veh,kw", [
    (1, 12, 3, {}), (2, 25, 4, {}), (3, 40, 5, dict(half_lo=300, half_hi=900)),
    (4, 40, 3, dict(half_lo=7200, half_hi=20000)), (5, 60, 6, {}), (6, 33, 2, dict(half_lo=20000, half_hi=43200)),
])
def test_algebra_equals_bruteforce(seed, stops, veh, kw):
    inst, node, _ = synth_vrptw(seed, stops, veh, **kw)
    rng = np.random.default_rng(seed)
    for state in (node, _random_state(inst, node, rng, 12)):
        assert O.route_state(inst, state)["feasible"]
        for fam in (0, 1, 2):
            db, fb = O.eval_bruteforce(inst, state, fam)
            df, ff = O.eval_fast(inst, state, fam)
            assert np.array_equal(fb, ff), f"feasibility flags differ, family {fam}"
            valid = fb != 2
            assert np.array_equal(db[valid], df[valid]), f"deltas differ, family {fam}"


def test_algebra_equals_bruteforce_with_distance_limit_and_asymmetry():
    inst, node, _ = synth_vrptw(7, 30, 4)
    rng = np.random.default_rng(7)
    D = inst.D.copy()
    D += rng.integers(0, 400, D.shape).astype(np.int32)  # asymmetric perturbation
    np.fill_diagonal(D, 0)
    T = inst.T + rng.integers(0, 40, D.shape).astype(np.int32)
    np.fill_diagonal(T, 0)
    base = O.Instance(D, T, inst.stops, inst.cap)
    rdist = O.route_state(base, node)["rdist"]
    lim = O.Instance(D, T, inst.stops, inst.cap, max_dist=int(rdist.max()) + 3000)
    assert O.route_state(lim, node)["feasible"]
    for fam in (0, 1, 2):
        db, fb = O.eval_bruteforce(lim, node, fam)
        df, ff = O.eval_fast(lim, node, fam)
        assert np.array_equal(fb, ff) and np.array_equal(db[fb != 2], df[fb != 2])
    assert (fb == 0).any() and (fb == 1).any()


def test_best_move_is_min_key_of_dumps_and_shards_partition():
    inst, node, _ = synth_vrptw(8, 50, 5)
    keys = []
    for fam in (0, 1, 2):
        d, f = O.eval_bruteforce(inst, node, fam)
        keys.append(O.best_from_dump(d, f, fam))
        assert O.best_move(inst, node, families=1 << fam) == keys[-1]
    assert O.best_move(inst, node) == min(keys)
    P = node.size
    cuts = [0, P // 3, P // 2, P]
    shard_best = [O.best_move(inst, node, shard=(cuts[i], cuts[i + 1]), threads=2) for i in range(3)]
    assert min(shard_best) == min(keys)


def test_candidate_counts_match_dumps():
    inst, node, _ = synth_vrptw(9, 37, 4)
    counts = O.count_candidates(node)
    for fam in (0, 1, 2):
        _, f = O.eval_bruteforce(inst, node, fam)
        assert counts[fam] == int((f != 2).sum())


def test_local_search_descends_and_stays_feasible():
    inst, node, _ = synth_vrptw(10, 45, 5)
    c0 = O.route_state(inst, node)["cost"]
    final, keys, cost = O.local_search(inst, node, threads=2)
    st = O.route_state(inst, final)
    assert st["feasible"] and st["cost"] == cost and cost < c0
    # replaying the keys reproduces the final state and the cost drops by exactly the sum of deltas
    replay = node.copy()
    total = 0
    for k in keys:
        d, fam, p, q = O.key_unpack(k)
        assert d < 0
        total += d
        replay = O.apply_move(replay, fam, p, q)
    assert np.array_equal(replay, final) and c0 + total == cost
    assert sorted(final[final != 0].tolist()) == list(range(1, inst.n))
    # local optimum: no improving feasible move is left
    d, _, _, _ = O.key_unpack(O.best_move(inst, final))
    assert d >= 0


def _route_minima_from_dumps(inst, node):
    """m[r] of SPEC §8.1 from the brute-force dumps: minimum key over the feasible candidates touching route r."""
    P, R = node.size, node.size - inst.n
    route = O.route_state(inst, node)["route"]
    m = np.full(R + 1, O.NO_MOVE, np.uint64)
    for fam in (0, 1, 2):
        d, f = O.eval_bruteforce(inst, node, fam)
        p, q = np.nonzero(f == 1)
        keys = ((d[p, q].astype(np.int64) + 0x80000000).astype(np.uint64) << np.uint64(32)) | \
            (np.uint64(fam) << np.uint64(30)) | (p.astype(np.uint64) << np.uint64(15)) | q.astype(np.uint64)
        for side in (p, q):
            np.minimum.at(m, route[side], keys)
    return m[:R], route


@pytest.mark.parametrize("seed,stops,veh", [(21, 40, 5), (22, 60, 8), (23, 30, 3)])
def test_batch_selection_is_pinned_to_bruteforce_and_moves_are_independent(seed, stops, veh):
    """SPEC §8.1: route minima equal the brute-force dumps'; the selected moves touch disjoint routes, contain the
    best move, and applying all of them changes the cost by exactly the sum of their deltas (each one's delta and
    feasibility survive the others)."""
    inst, node, _ = synth_vrptw(seed, stops, veh)
    rng = np.random.default_rng(seed)
    for state in (node, _random_state(inst, node, rng, 10)):
        m, keys = O.select_batch(inst, state, threads=2)
        want_m, route = _route_minima_from_dumps(inst, state)
        assert np.array_equal(m, want_m)
        best = O.best_move(inst, state)
        if O.key_unpack(best)[0] < 0:
            assert best in keys.tolist()
        else:
            assert keys.size == 0
        touched, total = [], 0
        for k in keys:
            d, fam, p, q = O.key_unpack(int(k))
            assert d < 0
            total += d
            touched += list({int(route[p]), int(route[q])})
        assert len(touched) == len(set(touched))
        first = O.select_batch(inst, state, rounds=1)[1]            # round 1 alone: ascending route[p], a prefix
        assert np.array_equal(first, keys[:first.size])
        assert [int(route[O.key_unpack(int(k))[2]]) for k in first] == sorted(int(route[O.key_unpack(int(k))[2]]) for k in first)
        after = O.apply_batch(inst, state, keys)
        st0, st1 = O.route_state(inst, state), O.route_state(inst, after)
        assert st1["feasible"] and st1["cost"] == st0["cost"] + total
        assert sorted(after[after != 0].tolist()) == list(range(1, inst.n))
        assert np.array_equal(O.apply_batch(inst, state, keys[:0]), state)


def test_batched_local_search_reaches_a_local_optimum_in_fewer_iterations():
    inst, node, _ = synth_vrptw(24, 80, 8)
    c0 = O.route_state(inst, node)["cost"]
    final, keys, sizes, cost = O.local_search_batched(inst, node, threads=2)
    _, seq_keys, seq_cost = O.local_search(inst, node, threads=2)
    st = O.route_state(inst, final)
    assert st["feasible"] and st["cost"] == cost and cost < c0
    assert sizes.sum() == keys.size and sizes.size < seq_keys.size and sizes.max() > 1
    assert c0 + sum(O.key_unpack(int(k))[0] for k in keys) == cost
    assert O.key_unpack(O.best_move(inst, final))[0] >= 0
    assert abs(cost - seq_cost) <= 0.15 * seq_cost          # a different local optimum of the same neighbourhood
    # max_iters stops early and resumes to the same end
    part, k1, s1, _ = O.local_search_batched(inst, node, max_iters=2, threads=2)
    assert s1.size == 2 and np.array_equal(k1, keys[:k1.size])
    rest, k2, _, c2 = O.local_search_batched(inst, part, threads=2)
    assert np.array_equal(rest, final) and c2 == cost and np.array_equal(np.concatenate([k1, k2]), keys)
