"""GPU: route state, move evaluation (every candidate's delta and feasibility flag), best-move keys and local
search through the C ABI, bit-exact against the oracle (whose algebra is itself pinned to brute force on the CPU)."""
import numpy as np
import pytest

from conftest import TS_CAR, synth_vrptw
from oracle import oracle as O

pytestmark = pytest.mark.gpu


def _load(engine, insts, states=None):
    """Problem holding the given oracle instances with the oracle's own matrices installed."""
    B, n = len(insts), insts[0][0].n
    pr = engine.problem(n, B)
    pr.set_coords(np.stack([r["lat"] for _, _, r in insts]), np.stack([r["lon"] for _, _, r in insts]))
    pr.build_matrix(10.0, TS_CAR)
    pr.set_stops(np.stack([i.stops for i, _, _ in insts]), [i.cap for i, _, _ in insts],
                 [i.max_dist for i, _, _ in insts])
    nodes = states if states is not None else [nd for _, nd, _ in insts]
    pr.set_routes([O.positions_to_trips(nd) for nd in nodes])
    return pr


def _walk(inst, node, seed, moves):
    rng = np.random.default_rng(seed)
    node = node.copy()
    for _ in range(moves):
        fam = int(rng.integers(0, 3))
        _, f = O.eval_fast(inst, node, fam)
        p, q = np.nonzero(f == 1)
        if p.size:
            i = int(rng.integers(0, p.size))
            node = O.apply_move(node, fam, int(p[i]), int(q[i]))
    return node


def test_device_matrix_equals_oracle_matrix(engine):
    inst, node, raw = synth_vrptw(21, 90, 6)
    with _load(engine, [(inst, node, raw)]) as pr:
        d, t = pr.get_matrix(0)
    assert np.array_equal(d, inst.D) and np.array_equal(t, inst.T)


def test_route_order_matrix_is_a_permutation_of_the_matrix(engine):
    """G[p][q] == M[node p][node q], both for the matrix rebuilt from coordinates and for a supplied one."""
    inst, node, raw = synth_vrptw(22, 130, 7)
    with _load(engine, [(inst, node, raw)]) as pr:
        P = node.size
        d, t = pr.get_position_matrix(0, P)
        assert np.array_equal(d, inst.D[np.ix_(node, node)]) and np.array_equal(t, inst.T[np.ix_(node, node)])
        pr.set_matrix(0, inst.D, inst.T)           # now the gather/permutation path
        pr.set_routes([O.positions_to_trips(node)])
        d2, t2 = pr.get_position_matrix(0, P)
        assert np.array_equal(d2, d) and np.array_equal(t2, t)


@pytest.mark.parametrize("seed,stops,veh,kw", [(1, 12, 3, {}), (2, 25, 4, {}), (3, 60, 5, dict(half_lo=300, half_hi=900)),
                                                (4, 100, 10, {}), (5, 150, 4, dict(half_lo=7200, half_hi=20000)),
                                                (6, 70, 2, dict(half_lo=20000, half_hi=43200))])
def test_route_state_and_every_candidate(engine, seed, stops, veh, kw):
    inst, node0, raw = synth_vrptw(seed, stops, veh, **kw)
    for node in (node0, _walk(inst, node0, seed, 15)):
        with _load(engine, [(inst, node0, raw)], [node]) as pr:
            st = pr.get_state(0)
            ref = O.route_state(inst, node)
            assert st["P"] == node.size and np.array_equal(st["node"], node)
            for k in ("route", "B", "Z", "rload", "rdist"):
                assert np.array_equal(st[k], ref[k]), k
            assert st["cost"] == ref["cost"] and st["feasible"] == ref["feasible"] is True
            best = []
            for fam in (0, 1, 2):
                d, f = pr.eval_dump(0, fam, node.size)
                dr, fr = O.eval_fast(inst, node, fam)
                assert np.array_equal(f, fr), f"feasibility flags differ in family {fam}"
                assert np.array_equal(d[fr != 2], dr[fr != 2]), f"deltas differ in family {fam}"
                best.append(O.best_from_dump(dr, fr, fam))
                assert int(pr.eval_moves(1 << fam)[0]) == best[-1]
            assert int(pr.eval_moves()[0]) == min(best) == O.best_move(inst, node)


def test_distance_limit_flags(engine):
    inst, node, raw = synth_vrptw(7, 80, 6)
    rdist = O.route_state(inst, node)["rdist"]
    lim = O.Instance(inst.D, inst.T, inst.stops, inst.cap, max_dist=int(rdist.max()) + 2000)
    with _load(engine, [(lim, node, raw)]) as pr:
        for fam in (0, 1, 2):
            d, f = pr.eval_dump(0, fam, node.size)
            dr, fr = O.eval_bruteforce(lim, node, fam)   # straight against brute force here
            assert np.array_equal(f, fr) and np.array_equal(d[fr != 2], dr[fr != 2])


def _perturbed(inst, seed, asymmetric):
    rng = np.random.default_rng(seed)
    bump = rng.integers(0, 300, inst.D.shape).astype(np.int32)
    if asymmetric:
        np.fill_diagonal(bump, 0)
        D, T = inst.D + bump, inst.T + bump // 10
    else:
        bump = np.triu(bump, 1)
        D, T = inst.D + bump + bump.T, inst.T + (bump + bump.T) // 10
    return O.Instance(D, T, inst.stops, inst.cap, inst.max_dist)


@pytest.mark.parametrize("asymmetric", [False, True], ids=["symmetric", "asymmetric"])
def test_caller_supplied_matrix(engine, asymmetric):
    """A matrix that is not great-circle (the ORS case: road distances, directional), installed through
    rt_problem_set_matrix: every candidate straight against brute force, then the search against the oracle."""
    inst, node, raw = synth_vrptw(8, 60, 5)
    alt = _perturbed(inst, 8, asymmetric)
    assert (alt.D != alt.D.T).any() == asymmetric
    # a feasible start on the new matrix: the oracle's window-aware greedy
    rc, trips = O.greedy_trips(alt.D, raw["demand"].astype(float), float(raw["cap"]), 9e12, T=alt.T, tw=raw["tw"])
    assert rc == 0
    node = O.build_positions(trips, alt.n)
    assert O.route_state(alt, node)["feasible"]
    with _load(engine, [(alt, node, raw)]) as pr:
        pr.set_matrix(0, alt.D, alt.T)
        d, t = pr.get_matrix(0)
        assert np.array_equal(d, alt.D) and np.array_equal(t, alt.T)
        g_trips, status = pr.greedy(raw["demand"][None].astype(float), [float(raw["cap"])], use_tw=True)
        assert status[0] == 0 and g_trips[0] == trips          # construction reads the matrix in the right direction
        pr.set_routes([trips])
        st = pr.get_state(0)
        ref = O.route_state(alt, node)
        assert st["cost"] == ref["cost"] and np.array_equal(st["B"], ref["B"]) and np.array_equal(st["Z"], ref["Z"])
        for fam in (0, 1, 2):
            dd, ff = pr.eval_dump(0, fam, node.size)
            dr, fr = O.eval_bruteforce(alt, node, fam)
            assert np.array_equal(ff, fr), f"flags, family {fam}"
            assert np.array_equal(dd[fr != 2], dr[fr != 2]), f"deltas, family {fam}"
        final, keys, cost = O.local_search(alt, node, threads=O.num_threads())
        moves, gcost, gkeys = pr.local_search(keys_capacity=4 * alt.n)
        assert moves[0] == len(keys) and np.array_equal(gkeys[0], keys) and gcost[0] == cost
        assert np.array_equal(pr.get_state(0)["node"], final)


def test_best_move_batched_and_sharded(engine):
    """A batch of instances with different route counts; the minimum over shards is the unsharded best move."""
    insts = [synth_vrptw(100 + b, 200, 5 + b) for b in range(6)]
    with _load(engine, insts) as pr:
        full = pr.eval_moves()
        for b, (inst, node, _) in enumerate(insts):
            assert int(full[b]) == O.best_move(inst, node, threads=O.num_threads())
        for count in (2, 3, 8):
            parts = np.stack([pr.eval_moves(shard=(i, count)) for i in range(count)])
            assert np.array_equal(parts.min(axis=0), full)


def test_plan_from_host_buffers(engine):
    """rt_problem_plan: host arrays in, best keys out, in one call — equals the step-by-step path and the oracle;
    a bad coordinate is reported (the range check runs on the device, off the host's critical path)."""
    from routest_b200 import EngineError
    insts = [synth_vrptw(400 + b, 150, 6) for b in range(4)]
    B, n = 4, 151
    with engine.problem(n, B) as pr:
        node = np.zeros((B, pr.pstride), np.int32)
        npos = np.zeros(B, np.int32)
        for b, (_, nd, _) in enumerate(insts):
            node[b, :nd.size] = nd
            npos[b] = nd.size
        lat = np.stack([r["lat"] for _, _, r in insts])
        lon = np.stack([r["lon"] for _, _, r in insts])
        stops = np.stack([i.stops for i, _, _ in insts])
        cap = np.array([i.cap for i, _, _ in insts], np.int32)
        md = np.full(B, 2 ** 62, np.int64)
        keys = pr.plan(lat, lon, stops, cap, md, node, npos, dist_scale=10.0, time_scale=TS_CAR)
        for b, (inst, nd, _) in enumerate(insts):
            assert int(keys[b]) == O.best_move(inst, nd)
        out = np.zeros(B, np.uint64)
        call = pr.plan_call(lat, lon, stops, cap, md, node, npos, out, dist_scale=10.0, time_scale=TS_CAR)
        assert np.array_equal(call(), keys) and call() is out
        # a directional matrix installed earlier does not leak into a pass that builds its own cells
        asym = insts[0][0].D + np.triu(np.full_like(insts[0][0].D, 50), 1)
        pr.set_matrix(0, asym, insts[0][0].T)
        assert np.array_equal(call(), keys)
        lat[2, 7] = 91.0
        with pytest.raises(EngineError) as ei:
            call()
        assert "coordinate" in str(ei.value)
        lat[2, 7] = insts[2][2]["lat"][7]
        assert np.array_equal(call(), keys)
        # what the kernels index with is checked before any of them runs: ids below n, exactly n-1 customers
        keep = node[1, 5]
        for bad in (n, -3, 0):
            node[1, 5] = bad
            with pytest.raises(EngineError) as ei:
                call()
            assert "position array" in str(ei.value)
        node[1, 5] = keep
        assert np.array_equal(call(), keys)


def test_plan_in_two_halves_with_two_problems_in_flight(engine):
    """rt_problem_plan_begin / rt_problem_plan_end: two problems of one engine used in turn (the upload of one request
    overlaps the kernels of the other) return what the synchronous call returns; misuse is refused."""
    from routest_b200 import EngineError
    B, n = 3, 121
    sets = [[synth_vrptw(500 + 10 * k + b, 120, 5) for b in range(B)] for k in range(4)]
    probs = [engine.problem(n, B) for _ in range(2)]
    try:
        bufs = []
        for pr in probs:
            bufs.append(dict(lat=np.zeros((B, n)), lon=np.zeros((B, n)), stops=np.zeros((B, n, 4), np.int32),
                             cap=np.zeros(B, np.int32), md=np.full(B, 2 ** 62, np.int64),
                             node=np.zeros((B, pr.pstride), np.int32), npos=np.zeros(B, np.int32), out=np.zeros(B, np.uint64)))
        calls = [pr.plan_pipeline_calls(b["lat"], b["lon"], b["stops"], b["cap"], b["md"], b["node"], b["npos"], b["out"],
                                        dist_scale=10.0, time_scale=TS_CAR) for pr, b in zip(probs, bufs)]

        def fill(b, insts):
            b["node"][:] = 0
            for i, (inst, nd, raw) in enumerate(insts):
                b["lat"][i], b["lon"][i], b["stops"][i], b["cap"][i] = raw["lat"], raw["lon"], inst.stops, inst.cap
                b["node"][i, :nd.size] = nd
                b["npos"][i] = nd.size

        want = [[O.best_move(inst, nd) for inst, nd, _ in insts] for insts in sets]
        got = []
        fill(bufs[0], sets[0])
        calls[0][0]()
        for k in range(1, len(sets)):
            fill(bufs[k % 2], sets[k])
            calls[k % 2][0]()                                   # request k goes out ...
            got.append(calls[(k - 1) % 2][1]().tolist())        # ... before request k-1 is collected
        got.append(calls[(len(sets) - 1) % 2][1]().tolist())
        assert got == want
        with pytest.raises(EngineError):
            calls[0][1]()                                       # nothing in flight
        calls[0][0]()
        with pytest.raises(EngineError):
            calls[0][0]()                                       # one pass per problem
        assert calls[0][1]().tolist() == want[2]                # buffers of problem 0 still hold request 2
    finally:
        for pr in probs:
            pr.close()


def test_replica_with_a_shard_keeps_only_its_rows_fresh(engine):
    """rt_problem_set_shard: the refresh rebuilds just the shard's rows of the route-order matrix (what each rank of
    a multi-GPU search does); every shard still finds its part of the best move, other shards are refused."""
    from routest_b200 import EngineError
    insts = [synth_vrptw(300 + b, 260, 9) for b in range(3)]
    with _load(engine, insts) as pr:
        full = pr.eval_moves()
        for count in (2, 5):
            parts = []
            for i in range(count):
                pr.set_shard(i, count)
                parts.append(pr.eval_moves(shard=(i, count)))
                with pytest.raises(EngineError):
                    pr.eval_moves(shard=((i + 1) % count, count))
            assert np.array_equal(np.stack(parts).min(axis=0), full)
        with pytest.raises(EngineError):
            pr.eval_dump(0, 1, insts[0][1].size)
        pr.set_shard(0, 1)
        assert np.array_equal(pr.eval_moves(), full)
        # state tables do not depend on the shard
        st = pr.get_state(1)
        ref = O.route_state(insts[1][0], insts[1][1])
        assert st["cost"] == ref["cost"] and np.array_equal(st["Z"], ref["Z"])


def test_apply_moves_matches_oracle(engine):
    inst, node, raw = synth_vrptw(9, 90, 6)
    with _load(engine, [(inst, node, raw)]) as pr:
        cur = node
        for fam in (1, 2, 0, 1, 0, 2, 1):
            key = int(pr.eval_moves(1 << fam)[0])
            assert key == O.best_move(inst, cur, families=1 << fam)
            if key == O.NO_MOVE:
                continue
            _, f, p, q = O.key_unpack(key)
            cur = O.apply_move(cur, f, p, q)
            pr.apply_moves(np.array([key], np.uint64))
            st = pr.get_state(0)
            assert np.array_equal(st["node"], cur) and st["feasible"]
            assert st["cost"] == O.route_state(inst, cur)["cost"]


def test_apply_moves_refuses_keys_that_are_not_moves(engine):
    """rt_problem_apply_moves: a key that is no move of the current routes leaves the instance as it is and is
    reported; 'no move' is accepted and does not block a later apply; the search flags are left alone."""
    from routest_b200 import EngineError
    inst, node, raw = synth_vrptw(19, 60, 5)
    P = node.size

    def key(fam, p, q):
        return (2 ** 31 << 32) | (fam << 30) | (p << 15) | q

    with _load(engine, [(inst, node, raw)]) as pr:
        pr.local_search()                                      # leaves every instance marked finished
        after = pr.get_state(0)["node"][:P].copy()
        sep = int(np.flatnonzero(after == 0)[1])               # an inner depot separator
        cust = int(np.flatnonzero(after != 0)[0])
        pr.apply_moves(np.array([O.NO_MOVE], np.uint64))
        for bad in (key(3, cust, cust + 1), key(2, 0, cust), key(2, cust, sep), key(1, sep, cust), key(1, cust, P - 1),
                    key(1, cust, cust), key(0, cust + 2, cust), key(0, cust, P - 1), key(2, cust, P + 3)):
            with pytest.raises(EngineError) as ei:
                pr.apply_moves(np.array([bad], np.uint64))
            assert "not a move" in str(ei.value)
            assert np.array_equal(pr.get_state(0)["node"][:P], after)
        a, b = [int(x) for x in np.flatnonzero(after != 0)[[1, 4]]]
        pr.apply_moves(np.array([key(2, a, b)], np.uint64))     # a real swap still goes through
        want = O.apply_move(after, 2, a, b)
        assert np.array_equal(pr.get_state(0)["node"][:P], want)


@pytest.mark.parametrize("full", [False, True], ids=["incremental", "full"])
@pytest.mark.parametrize("seed,stops,veh", [(10, 45, 5), (11, 100, 10), (12, 300, 12)])
def test_local_search_moves_and_final_routes_bit_exact(engine, seed, stops, veh, full, monkeypatch):
    """Both search drivers — the one that re-scores only what a move touched (per-route-pair table) and the one that
    re-scores everything on the route-order matrix — apply the oracle's exact sequence of keys.  (By default the
    incremental driver only takes over above ~1,500 positions; the environment switch forces it here.)"""
    monkeypatch.setenv("ROUTEST_LS_INCREMENTAL_FROM", "0")
    inst, node, raw = synth_vrptw(seed, stops, veh)
    final, keys, cost = O.local_search(inst, node, threads=O.num_threads())
    with _load(engine, [(inst, node, raw)]) as pr:
        moves, gcost, gkeys = pr.local_search(keys_capacity=4 * inst.n, full=full)
        st = pr.get_state(0)
        trips = pr.get_routes()[0]
    assert moves[0] == len(keys) and np.array_equal(gkeys[0], keys)
    assert gcost[0] == cost and st["cost"] == cost and st["feasible"]
    assert np.array_equal(st["node"], final)
    assert trips == O.positions_to_trips(final)


@pytest.mark.parametrize("full", [False, True], ids=["incremental", "full"])
def test_local_search_batch_and_max_moves(engine, full, monkeypatch):
    monkeypatch.setenv("ROUTEST_LS_INCREMENTAL_FROM", "0")
    insts = [synth_vrptw(200 + b, 80, 6) for b in range(4)]
    refs = [O.local_search(inst, node, max_moves=25) for inst, node, _ in insts]
    with _load(engine, insts) as pr:
        moves, cost, keys = pr.local_search(max_moves=25, keys_capacity=64, full=full)
        for b in range(4):
            assert moves[b] == len(refs[b][1]) and np.array_equal(keys[b], refs[b][1]) and cost[b] == refs[b][2]
            assert np.array_equal(pr.get_state(b)["node"], refs[b][0])


@pytest.mark.parametrize("seed,stops,veh", [(10, 45, 5), (11, 100, 10), (12, 300, 12), (14, 500, 40), (16, 640, 300)])
def test_batched_local_search_applies_the_oracles_batches(engine, seed, stops, veh):
    """SPEC §8.1 (rt_problem_local_search_batched): every iteration's set of independent moves, key for key, the
    final routes and the cost equal the oracle's; the end state is a local optimum of the full neighbourhood.  Up to
    128 routes the selection and the apply step are one kernel, beyond that one kernel per step (last case)."""
    inst, node, raw = synth_vrptw(seed, stops, veh)
    assert (node.size - inst.n > 128) == (veh > 128)
    final, keys, sizes, cost = O.local_search_batched(inst, node, threads=O.num_threads())
    with _load(engine, [(inst, node, raw)]) as pr:
        moves, gcost, gkeys = pr.local_search(keys_capacity=4 * inst.n, batched=True)
        st = pr.get_state(0)
        best_after = int(pr.eval_moves()[0])
    assert sizes.max() > 1 and sizes.size < keys.size
    assert moves[0] == len(keys) and np.array_equal(gkeys[0], keys)
    assert gcost[0] == cost and st["cost"] == cost and st["feasible"]
    assert np.array_equal(st["node"], final)
    assert O.key_unpack(best_after)[0] >= 0


@pytest.mark.parametrize("seed,stops,veh", [(1, 12, 3), (2, 6, 1), (3, 2, 1), (4, 1, 1)])
def test_batched_local_search_on_tiny_instances(engine, seed, stops, veh):
    """Edge sizes: a handful of stops, a single route, nothing to improve at all."""
    inst, node, raw = synth_vrptw(seed, stops, veh)
    final, keys, sizes, cost = O.local_search_batched(inst, node)
    with _load(engine, [(inst, node, raw)]) as pr:
        moves, gcost, gkeys = pr.local_search(keys_capacity=64, batched=True)
        assert moves[0] == keys.size and np.array_equal(gkeys[0], keys) and gcost[0] == cost
        assert np.array_equal(pr.get_state(0)["node"], final)
        moves, gcost, _ = pr.local_search(batched=True)          # already a local optimum: nothing happens
        assert moves[0] == 0 and gcost[0] == cost


def test_batched_local_search_batch_of_instances_max_iterations_and_supplied_matrix(engine):
    insts = [synth_vrptw(210 + b, 90, 5 + b) for b in range(4)]          # same size, different route counts
    refs = [O.local_search_batched(inst, node, max_iters=3) for inst, node, _ in insts]
    ends = [O.local_search_batched(inst, node) for inst, node, _ in insts]
    with _load(engine, insts) as pr:
        moves, cost, keys = pr.local_search(max_moves=3, keys_capacity=256, batched=True)
        for b in range(4):
            assert moves[b] == len(refs[b][1]) and np.array_equal(keys[b], refs[b][1]) and cost[b] == refs[b][3]
            assert np.array_equal(pr.get_state(b)["node"], refs[b][0])
        moves, cost, _ = pr.local_search(batched=True)                   # resumes to the oracle's end state
        for b in range(4):
            assert cost[b] == ends[b][3] and np.array_equal(pr.get_state(b)["node"], ends[b][0])
    # a caller-supplied asymmetric matrix goes through the same path
    inst, node, raw = insts[0]
    rng = np.random.default_rng(5)
    D = inst.D + rng.integers(0, 300, inst.D.shape).astype(np.int32)
    T = inst.T + rng.integers(0, 30, inst.T.shape).astype(np.int32)
    np.fill_diagonal(D, 0)
    np.fill_diagonal(T, 0)
    asym = O.Instance(D, T, inst.stops, inst.cap)
    if O.route_state(asym, node)["feasible"]:
        want = O.local_search_batched(asym, node)
        with _load(engine, [(inst, node, raw)]) as pr:
            pr.set_matrix(0, D, T)
            moves, cost, keys = pr.local_search(keys_capacity=1024, batched=True)
            assert np.array_equal(keys[0], want[1]) and cost[0] == want[3]
            assert np.array_equal(pr.get_state(0)["node"], want[0])


def test_batched_search_through_sharded_table_entry_points(engine):
    """The _dev steps of a multi-GPU batched search (rt_problem_eval_table_dev / rt_problem_apply_batch_dev) with the
    int64 MIN all-reduce replaced by torch.minimum over the shards' tables: the oracle's batches, for two instances."""
    import torch
    insts = [synth_vrptw(15 + b, 140, 7 + b) for b in range(2)]
    ends = [O.local_search_batched(inst, node, threads=O.num_threads()) for inst, node, _ in insts]
    shards = 3
    with _load(engine, insts) as pr:
        s = torch.cuda.current_stream().cuda_stream
        tabs = [torch.empty(pr.table_entries(), dtype=torch.int64, device="cuda") for _ in range(shards)]
        for it in range(max(e[2].size for e in ends) + 3):
            for i in range(shards):
                pr.eval_table_dev(tabs[i].data_ptr(), shard=(i, shards), stream=s)
            combined = torch.stack(tabs).min(dim=0).values.contiguous()
            pr.apply_batch_dev(combined.data_ptr(), stream=s)
        done, moves, gcost = pr.search_status()
        for b in range(2):
            assert done[b] == 1 and moves[b] == ends[b][1].size and gcost[b] == ends[b][3]
            assert np.array_equal(pr.get_state(b)["node"], ends[b][0])


def test_multi_shard_search_through_dev_entry_points(engine):
    """The two _dev steps a multi-GPU search is made of, with the NCCL MIN replaced by torch.minimum over the
    shards' key tensors (one GPU here): same applied moves and final routes as the oracle."""
    import torch
    inst, node, raw = synth_vrptw(13, 120, 8)
    final, keys, cost = O.local_search(inst, node, threads=O.num_threads())
    shards = 4
    with _load(engine, [(inst, node, raw)]) as pr:
        s = torch.cuda.current_stream().cuda_stream
        bufs = [torch.empty(1, dtype=torch.int64, device="cuda") for _ in range(shards)]
        for it in range(len(keys) + 5):
            for i in range(shards):
                pr.eval_moves_dev(bufs[i].data_ptr(), shard=(i, shards), stream=s)
            best = torch.stack(bufs).min(dim=0).values.contiguous()
            pr.apply_moves_dev(best.data_ptr(), improving_only=True, stream=s)
        done, moves, gcost = pr.search_status()
        st = pr.get_state(0)
    assert done[0] == 1 and moves[0] == len(keys) and gcost[0] == cost
    assert np.array_equal(st["node"], final)


def _tiny_instance(stops, tw_due=1000):
    rng = np.random.default_rng(stops)
    n = stops + 1
    lat = np.concatenate([[14.58], 14.58 + 0.01 * rng.uniform(-1, 1, stops)])
    lon = np.concatenate([[121.04], 121.04 + 0.01 * rng.uniform(-1, 1, stops)])
    m = O.matrix_rows(O.unit_vectors(lat, lon), 0, n, 10.0, TS_CAR)
    st = np.zeros((n, 4), np.int32)
    st[:, 1] = tw_due
    st[1:, 3] = 1
    return O.Instance(m["dist"], m["time"], st, 5), dict(lat=lat, lon=lon)


@pytest.mark.parametrize("trips", [[[1]], [[1], [2]], [[2, 1]], [[1, 2], [3]], [[3], [1], [2]]], ids=str)
def test_tiny_instances_and_routes_that_become_empty(engine, trips):
    """One to three stops, one route each or together: every candidate against brute force, then the search
    (which may empty a route: two adjacent separators) against the oracle."""
    stops = sum(len(t) for t in trips)
    inst, raw = _tiny_instance(stops)
    node = O.build_positions(trips, inst.n)
    with _load(engine, [(inst, node, raw)]) as pr:
        for fam in (0, 1, 2):
            d, f = pr.eval_dump(0, fam, node.size)
            dr, fr = O.eval_bruteforce(inst, node, fam)
            assert np.array_equal(f, fr) and np.array_equal(d[fr != 2], dr[fr != 2])
        assert int(pr.eval_moves()[0]) == O.best_move(inst, node)
        final, keys, cost = O.local_search(inst, node)
        moves, gcost, gkeys = pr.local_search(keys_capacity=16)
        assert moves[0] == len(keys) and np.array_equal(gkeys[0], keys) and gcost[0] == cost
        st = pr.get_state(0)
        assert np.array_equal(st["node"], final) and st["feasible"]
        # after the search routes may be empty; evaluation on that state still matches
        for fam in (0, 1, 2):
            d, f = pr.eval_dump(0, fam, final.size)
            dr, fr = O.eval_bruteforce(inst, final, fam)
            assert np.array_equal(f, fr) and np.array_equal(d[fr != 2], dr[fr != 2])
        assert pr.get_routes()[0] == O.positions_to_trips(final)


def test_local_search_with_zero_moves_and_bad_routes(engine):
    from routest_b200 import EngineError
    inst, node, raw = synth_vrptw(14, 30, 3)
    with _load(engine, [(inst, node, raw)]) as pr:
        moves, cost, _ = pr.local_search(max_moves=0)
        assert moves[0] == 0 and cost[0] == O.route_state(inst, node)["cost"]
        with pytest.raises(EngineError):
            pr.set_routes([[[1, 2, 3]]])                      # not all customers
        with pytest.raises(EngineError):
            pr.set_routes([[[1, 1] + list(range(3, 31))]])    # a customer twice


def test_peer_exchange_entry_points_single_replica(engine):
    """The exchange path a multi-GPU search uses instead of NCCL (slot buffer, publish with system-scope atomics,
    bounded wait, re-arm), here with one replica: same moves and routes as the oracle.  The multi-process form is
    exercised by tools/dist_search_check.py --exchange peer on 2 and 8 GPUs."""
    inst, node, raw = synth_vrptw(15, 90, 6)
    final, keys, cost = O.local_search(inst, node, threads=O.num_threads())
    with _load(engine, [(inst, node, raw)]) as pr:
        handle = pr.exchange_export()
        assert len(handle) == 64
        pr.exchange_connect([handle], 0, 1)
        for _ in range(len(keys) + 3):
            pr.eval_moves_xchg_dev()
            pr.apply_moves_xchg_dev(True)
        done, moves, gcost = pr.search_status()
        assert done[0] == 1 and moves[0] == len(keys) and gcost[0] == cost
        assert np.array_equal(pr.get_state(0)["node"], final)


def test_table_exchange_entry_points_single_replica(engine):
    """The batched search's table through the peer exchange (all-gather into the replicas' buffers + local minimum),
    here with one replica and many iterations through both slots: the oracle's batches, key for key."""
    import torch
    from routest_b200.dist import DistributedSearch
    inst, node, raw = synth_vrptw(16, 180, 7)
    final, okeys, sizes, ocost = O.local_search_batched(inst, node, threads=O.num_threads())
    for graph in (False, True):
        with _load(engine, [(inst, node, raw)]) as pr:
            search = DistributedSearch(pr, 1, device="cuda", check_every=3, exchange="peer", batched=True, graph=graph)
            moves, cost = search.run()
            assert moves[0] == okeys.size and cost[0] == ocost
            assert np.array_equal(pr.get_state(0)["node"], final)
            assert (search._graph is not None) == graph


def test_plan_upload_dev_scores_the_replica_shard(engine):
    """rt_problem_plan_upload_dev: host inputs in, the keys of THIS replica's shard left on the device in signed-order
    form; the minimum over the shards (what the NCCL MIN computes across ranks) is the oracle's best move."""
    import torch
    inst, node, raw = synth_vrptw(77, 700, 8)
    s = torch.cuda.current_stream().cuda_stream
    keys = torch.zeros(1, dtype=torch.int64, device="cuda")
    found = []
    for world in (1, 3):
        for rank in range(world):
            with engine.problem(inst.n, 1) as pr:
                pr.set_shard(rank, world)
                nd = np.zeros((1, pr.pstride), np.int32)
                nd[0, :node.size] = node
                call = pr.plan_upload_call(np.ascontiguousarray(raw["lat"][None]), np.ascontiguousarray(raw["lon"][None]),
                                           np.ascontiguousarray(inst.stops[None]), np.array([inst.cap], np.int32),
                                           np.full(1, 2 ** 62, np.int64), nd, np.array([node.size], np.int32),
                                           keys.data_ptr(), dist_scale=10.0, time_scale=TS_CAR, stream=s)
                call()
                call()   # a second request on the same problem: nothing left over from the first
                torch.cuda.synchronize()
                found.append((world, (int(keys.item()) ^ -(2 ** 63)) & (2 ** 64 - 1)))
    want = O.best_move(inst, node)
    assert found[0] == (1, want)
    assert min(k for w, k in found if w == 3) == want
    with engine.problem(inst.n, 1) as pr:   # a coordinate out of range is refused before anything is enqueued
        nd = np.zeros((1, pr.pstride), np.int32)
        nd[0, :node.size] = node
        lat = np.ascontiguousarray(raw["lat"][None]).copy()
        lat[0, 5] = 91.0
        call = pr.plan_upload_call(lat, np.ascontiguousarray(raw["lon"][None]), np.ascontiguousarray(inst.stops[None]),
                                   np.array([inst.cap], np.int32), np.full(1, 2 ** 62, np.int64), nd,
                                   np.array([node.size], np.int32), keys.data_ptr(), dist_scale=10.0, time_scale=TS_CAR, stream=s)
        from routest_b200 import EngineError
        with pytest.raises(EngineError):
            call()


def test_state_tables_refresh_alone_follows_a_route_change(engine):
    """rt_problem_refresh_dev(what=2) on its own after the routes changed: the tables re-derive their unit vectors
    themselves (they do not depend on an earlier what=1 refresh)."""
    import torch
    inst, node, raw = synth_vrptw(78, 300, 4)
    with _load(engine, [(inst, node, raw)]) as pr:
        k = int(pr.eval_moves()[0])
        keys = torch.from_numpy((np.array([k], np.uint64) ^ np.uint64(1 << 63)).view(np.int64)).cuda()
        s = torch.cuda.current_stream().cuda_stream
        pr.apply_moves_dev(keys.data_ptr(), improving_only=True, stream=s)   # routes change (and are refreshed)
        torch.cuda.synchronize()
        d, f, p, q = O.key_unpack(k)
        node2 = O.apply_move(node, f, p, q)
        pr.refresh_dev(2, s)            # tables only
        torch.cuda.synchronize()
        st, ref = pr.get_state(0), O.route_state(inst, node2)
        assert st["cost"] == ref["cost"] and np.array_equal(st["B"], ref["B"]) and np.array_equal(st["Z"], ref["Z"])
        pr.refresh_dev(3, s)
        assert int(pr.eval_moves()[0]) == O.best_move(inst, node2)


def test_measured_pipe_peaks_are_plausible(engine):
    """rt_selftest_pipe_peak: the ALU pipe issues one warp instruction every two clocks per scheduler, the issue limit
    is one per clock — the measured rates must sit just under 148 SMs x 4 schedulers x clock x {1/2, 1}."""
    alu, _ = engine.pipe_peak(0, 20000)
    both, _ = engine.pipe_peak(2, 20000)
    lo, hi = 148 * 4 * 1.0e9, 148 * 4 * 2.1e9
    assert 0.5 * lo * 0.8 < alu < 0.5 * hi * 1.02, alu
    assert lo * 0.8 < both < hi * 1.02, both
    assert 1.6 < both / alu < 2.1


@pytest.mark.parametrize("fused", [1, 0], ids=["cells-from-vectors", "route-order-matrix"])
@pytest.mark.parametrize("kernel", [1, 2], ids=["thread-per-anchor", "warp-per-anchor"])
@pytest.mark.parametrize("seed,stops,veh,kw", [(11, 40, 6, {}), (12, 150, 4, dict(half_lo=7200, half_hi=20000)),
                                                (13, 260, 3, dict(half_lo=20000, half_hi=43200))])
def test_kernel_variants_agree_with_the_oracle(engine, fused, kernel, seed, stops, veh, kw):
    """Every variant of the move kernels scores every candidate like the oracle: the same-route kernels (sweep per
    thread / segment scans per warp; rt_debug_set "intra_kernel") on routes of 7, ~40 and ~90 positions — no carry,
    one carry, several carries between the warp's 32-position chunks — and both sources of matrix cells ("fused_eval":
    rebuilt from the unit vectors inside the kernels / read from the route-order matrix); same table-mode batches and
    searches too."""
    inst, node0, raw = synth_vrptw(seed, stops, veh, **kw)
    engine.debug_set("intra_kernel", kernel)
    engine.debug_set("fused_eval", fused)
    try:
        for node in (node0, _walk(inst, node0, seed, 10)):
            with _load(engine, [(inst, node0, raw)], [node]) as pr:
                for fam in (0, 1, 2):
                    d, f = pr.eval_dump(0, fam, node.size)
                    dr, fr = O.eval_fast(inst, node, fam)
                    assert np.array_equal(f, fr), f"feasibility flags differ in family {fam}"
                    assert np.array_equal(d[fr != 2], dr[fr != 2]), f"deltas differ in family {fam}"
                assert int(pr.eval_moves()[0]) == O.best_move(inst, node)
                parts = [int(pr.eval_moves(shard=(i, 3))[0]) for i in range(3)]
                assert min(parts) == O.best_move(inst, node)
        with _load(engine, [(inst, node0, raw)]) as pr:
            bmoves, bcost, bkeys = pr.local_search(max_moves=3, keys_capacity=512, batched=True)   # table mode
            bfinal, okeys, _, ocost = O.local_search_batched(inst, node0, max_iters=3)
            assert np.array_equal(bkeys[0], okeys) and bcost[0] == ocost
            moves, cost, keys = pr.local_search(max_moves=12, keys_capacity=16, full=True)
            final, okeys2, ocost2 = O.local_search(inst, bfinal, max_moves=12)
            assert np.array_equal(keys[0], okeys2) and cost[0] == ocost2
            assert np.array_equal(pr.get_state(0)["node"], final)
            P = node0.size   # the route-order matrix itself is still there for whoever asks
            d, t = pr.get_position_matrix(0, pr.get_state(0)["P"])
            nd = pr.get_state(0)["node"]
            assert np.array_equal(d, inst.D[np.ix_(nd, nd)]) and np.array_equal(t, inst.T[np.ix_(nd, nd)])
    finally:
        engine.debug_set("intra_kernel", 0)
        engine.debug_set("fused_eval", 1)


@pytest.mark.parametrize("persistent,state_fused_max,apply_prep,small_strips",
                         [(2, -1, 1, 1), (2, -1, 1, 0), (0, -1, 1, 1), (0, -1, 0, 1), (0, 4096, 1, 0), (0, 4096, 0, 0)],
                         ids=["one-launch", "one-launch-32-rows", "merged-apply", "separate-apply", "one-block-state",
                              "round-1-chain"])
@pytest.mark.parametrize("seed,stops,veh,kw", [(21, 12, 1, {}), (22, 150, 4, dict(half_lo=7200, half_hi=20000)),
                                                (23, 420, 40, {})])
def test_search_launch_variants_agree_with_the_oracle(engine, persistent, state_fused_max, apply_prep, small_strips, seed,
                                                      stops, veh, kw):
    """The one-move search reaches the oracle's keys, cost and routes whichever way an iteration is cut into launches:
    the whole search as one cooperative launch (rt_debug_set "persistent_search" = 2: wherever it can run, not only
    at the small sizes where it is the default); the move + route scan + per-route
    tables in one launch ("apply_prep"; more than 32 routes leave the tables to their own kernel: the 40-vehicle
    case); the one-block state kernel or one warp per route ("state_fused_max"); short strips for small grids
    ("small_strips").  A search cut short by max_moves and then resumed ends where the uninterrupted one does.
    SPEC.md §8; no reference counterpart."""
    inst, node0, raw = synth_vrptw(seed, stops, veh, **kw)
    engine.debug_set("persistent_search", persistent)
    engine.debug_set("state_fused_max", state_fused_max)
    engine.debug_set("apply_prep", apply_prep)
    engine.debug_set("small_strips", small_strips)
    try:
        with _load(engine, [(inst, node0, raw)]) as pr:
            assert int(pr.eval_moves()[0]) == O.best_move(inst, node0)
            moves, cost, keys = pr.local_search(max_moves=40, keys_capacity=64, full=True)
            final, okeys, ocost = O.local_search(inst, node0, max_moves=40)
            assert np.array_equal(keys[0], okeys) and cost[0] == ocost
            assert np.array_equal(pr.get_state(0)["node"], final)
            assert int(pr.eval_moves()[0]) == O.best_move(inst, final)
            st = pr.get_state(0)
            ref = O.route_state(inst, final)
            for k in ("route", "B", "Z", "rload", "rdist"):
                assert np.array_equal(st[k], ref[k]), k
            moves2, cost2, _ = pr.local_search(full=True)                      # on to the local optimum
            final2, okeys2, ocost2 = O.local_search(inst, final)
            assert moves2[0] == len(okeys2) and cost2[0] == ocost2
            assert np.array_equal(pr.get_state(0)["node"], final2)
    finally:
        engine.debug_set("persistent_search", 1)
        engine.debug_set("state_fused_max", -1)
        engine.debug_set("apply_prep", 1)
        engine.debug_set("small_strips", 1)


def test_one_launch_search_on_a_batch(engine):
    """The cooperative search kernel with several instances of different route counts in one problem: each stops at
    its own local optimum (done[] per instance), the others go on."""
    insts = [synth_vrptw(300 + b, 100, 3 + b) for b in range(5)]
    n = insts[0][0].n
    engine.debug_set("persistent_search", 2)
    try:
        with _load(engine, insts) as pr:
            moves, cost, keys = pr.local_search(keys_capacity=4 * n, full=True)
            for b, (inst, node, _) in enumerate(insts):
                final, okeys, ocost = O.local_search(inst, node, threads=O.num_threads())
                assert moves[b] == len(okeys) and np.array_equal(keys[b], okeys) and cost[b] == ocost
                assert np.array_equal(pr.get_state(b)["node"], final)
    finally:
        engine.debug_set("persistent_search", 1)


@pytest.mark.parametrize("seed,stops,veh", [(31, 80, 6), (32, 300, 12)])
def test_binding_distance_limit_through_search(engine, seed, stops, veh):
    """A distance limit that binds (the median route length: half of the routes are already over it and accept
    nothing, the others have little room): best move, one-move search and batched
    iterations against the oracle.  Without a limit the tile kernel runs a form with the slack tests compiled out
    (every other test of this file); here they decide which candidates are feasible."""
    inst, node, raw = synth_vrptw(seed, stops, veh)
    rdist = O.route_state(inst, node)["rdist"]
    lim = O.Instance(inst.D, inst.T, inst.stops, inst.cap, max_dist=int(np.median(rdist)) + 200)
    lim_key = O.best_move(lim, node)
    assert lim_key != O.best_move(inst, node)   # the limit decides the very first move, or this test tests nothing
    with _load(engine, [(lim, node, raw)]) as pr:
        assert int(pr.eval_moves()[0]) == lim_key
        parts = [int(pr.eval_moves(shard=(i, 3))[0]) for i in range(3)]
        assert min(parts) == lim_key
        moves, cost, keys = pr.local_search(max_moves=30, keys_capacity=64, full=True)
        final, okeys, ocost = O.local_search(lim, node, max_moves=30)
        assert np.array_equal(keys[0], okeys) and cost[0] == ocost
        assert np.array_equal(pr.get_state(0)["node"], final)
    with _load(engine, [(lim, node, raw)]) as pr:
        bmoves, bcost, bkeys = pr.local_search(max_moves=4, keys_capacity=1024, batched=True)
        bfinal, obkeys, _, obcost = O.local_search_batched(lim, node, max_iters=4)
        assert np.array_equal(bkeys[0], obkeys) and bcost[0] == obcost


@pytest.mark.parametrize("seed,stops,veh,kw", [(41, 60, 5, dict(half_lo=300, half_hi=900)), (42, 150, 4, {})])
def test_unlimited_capacity_meets_the_validity_sentinels(engine, seed, stops, veh, kw):
    """Capacity 2^31 - 1: every capacity slack sits at its clamp (2^30), the largest value the tile kernel ever
    compares a demand with — the stand-in demand of a stop that may not be moved has to stay above it.  Tight windows
    make many stops immovable (removing them would break their route's schedule)."""
    inst, node, raw = synth_vrptw(seed, stops, veh, **kw)
    big = O.Instance(inst.D, inst.T, inst.stops, 2 ** 31 - 1, inst.max_dist)
    with _load(engine, [(big, node, raw)]) as pr:
        assert int(pr.eval_moves()[0]) == O.best_move(big, node)
        moves, cost, keys = pr.local_search(max_moves=60, keys_capacity=64, full=True)
        final, okeys, ocost = O.local_search(big, node, max_moves=60)
        assert np.array_equal(keys[0], okeys) and cost[0] == ocost
        assert np.array_equal(pr.get_state(0)["node"], final)
    with _load(engine, [(big, node, raw)]) as pr:
        bmoves, bcost, bkeys = pr.local_search(max_moves=4, keys_capacity=1024, batched=True)
        _, obkeys, _, obcost = O.local_search_batched(big, node, max_iters=4)
        assert np.array_equal(bkeys[0], obkeys) and bcost[0] == obcost
