"""CPU, world_size 2, gloo: the host side of the multi-GPU path — range sharding and the replicated search loop with
its MIN all-reduce of signed-order keys.  The GPU steps are replaced by a stand-in backed by the oracle, so what is
tested is routest_b200/dist.py itself: both ranks must walk the oracle's exact move sequence."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

from conftest import ROOT


def test_shard_range_partitions():
    from routest_b200.dist import shard_range
    for total in (0, 1, 7, 64, 1001):
        for world in (1, 2, 3, 8):
            cuts = [shard_range(total, r, world) for r in range(world)]
            assert cuts[0][0] == 0 and cuts[-1][1] == total
            assert all(cuts[i][1] == cuts[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in cuts]
            assert max(sizes) - min(sizes) <= 1


class OracleProblem:
    """Stand-in for routest_b200.Problem on the CPU: same two stream-ordered steps, evaluated by the oracle."""

    def __init__(self, inst, node):
        from oracle import oracle as O
        self.O, self.inst, self.node = O, inst, node.copy()
        self.done, self.moves = 0, 0

    def _tensor(self, ptr):
        import ctypes
        return (ctypes.c_int64 * 1).from_address(ptr)

    def eval_moves_dev(self, keys_ptr, families, shard, stream):
        P = self.node.size
        i, cnt = shard
        k = self.O.best_move(self.inst, self.node, families, shard=(P * i // cnt, P * (i + 1) // cnt))
        self._tensor(keys_ptr)[0] = (k ^ (1 << 63)) - (1 << 64) if (k ^ (1 << 63)) >= (1 << 63) else (k ^ (1 << 63))

    def apply_moves_dev(self, keys_ptr, improving_only, stream):
        if self.done:
            return
        signed = self._tensor(keys_ptr)[0]
        k = (signed & (2 ** 64 - 1)) ^ (1 << 63)
        d, f, p, q = self.O.key_unpack(k)
        if k == self.O.NO_MOVE or (improving_only and d >= 0):
            self.done = 1
            return
        self.node = self.O.apply_move(self.node, f, p, q)
        self.moves += 1

    def search_status(self):
        return [self.done], [self.moves], [self.O.route_state(self.inst, self.node)["cost"]]


class OracleBatchProblem(OracleProblem):
    """Stand-in for the batched steps (SPEC §8.1): the shard's table of best keys per ordered route pair comes from
    the oracle's candidate dumps, the selection from the MIN-reduced table is restated here in numpy."""

    NONE = np.uint64(2 ** 64 - 1)
    FLIP = np.uint64(1 << 63)

    def table_entries(self):
        R = self.node.size - self.inst.n
        return R * R + R

    def _view(self, ptr):
        import ctypes
        n = self.table_entries()
        return np.frombuffer((ctypes.c_uint64 * n).from_address(ptr), dtype=np.uint64)

    def eval_table_dev(self, table_ptr, families, shard, stream):
        O, P, R = self.O, self.node.size, self.node.size - self.inst.n
        route = O.route_state(self.inst, self.node)["route"]
        i, cnt = shard
        p0, p1 = P * i // cnt, P * (i + 1) // cnt
        inter = np.full((R, R), self.NONE, np.uint64)
        intra = np.full(R, self.NONE, np.uint64)
        for fam in (0, 1, 2):
            if not families & (1 << fam):
                continue
            d, f = O.eval_fast(self.inst, self.node, fam)
            for p, q in zip(*np.nonzero(f == 1)):
                if not p0 <= p < p1:
                    continue
                k = np.uint64(((int(d[p, q]) + 2 ** 31) << 32) | (fam << 30) | (int(p) << 15) | int(q))
                ra, rb = route[p], route[q]
                if ra == rb:
                    intra[ra] = min(intra[ra], k)
                else:
                    inter[ra, rb] = min(inter[ra, rb], k)
        self._view(table_ptr)[:] = np.concatenate([inter.ravel(), intra]) ^ self.FLIP

    def apply_batch_dev(self, table_ptr, stream):
        if self.done:
            return
        O, R = self.O, self.node.size - self.inst.n
        route = O.route_state(self.inst, self.node)["route"]
        t = self._view(table_ptr) ^ self.FLIP
        inter, intra = t[:R * R].reshape(R, R), t[R * R:]
        taken, batch = np.zeros(R, bool), []
        for _ in range(O.BATCH_ROUNDS):
            m = np.full(R, self.NONE, np.uint64)
            for ra in range(R):
                if taken[ra]:
                    continue
                m[ra] = min(m[ra], intra[ra])
                for rb in range(R):
                    if not taken[rb]:
                        m[ra], m[rb] = min(m[ra], inter[ra, rb]), min(m[rb], inter[ra, rb])
            sel = []
            for r in range(R):
                k = int(m[r])
                if k == O.NO_MOVE or O.key_unpack(k)[0] >= 0:
                    continue
                _, _, pp, qq = O.key_unpack(k)
                if route[pp] == r and int(m[route[qq]]) == k:
                    sel.append(k)
            if not sel:
                break
            for k in sel:
                taken[route[O.key_unpack(k)[2]]] = taken[route[O.key_unpack(k)[3]]] = True
            batch += sel
        if not batch:
            self.done = 1
            return
        self.node = O.apply_batch(self.inst, self.node, np.array(batch, np.uint64))
        self.moves += len(batch)


def _worker_batched(rank, world, port, out):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    import torch.distributed as dist
    from conftest import synth_vrptw
    from routest_b200.dist import DistributedSearch
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        inst, node, _ = synth_vrptw(32, 50, 5)
        prob = OracleBatchProblem(inst, node)
        search = DistributedSearch(prob, 1, device="cpu", check_every=4, batched=True)
        moves, cost = search.run()
        out[rank] = (int(moves[0]), int(cost[0]), prob.node.tolist())
    finally:
        dist.destroy_process_group()


def _worker(rank, world, port, out):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    import torch.distributed as dist
    from conftest import synth_vrptw
    from routest_b200.dist import DistributedSearch
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        inst, node, _ = synth_vrptw(31, 60, 5)
        prob = OracleProblem(inst, node)
        search = DistributedSearch(prob, 1, device="cpu", check_every=8)
        moves, cost = search.run()
        out[rank] = (int(moves[0]), int(cost[0]), prob.node.tolist())
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_replicated_search_under_gloo_matches_oracle():
    from oracle import oracle as O
    from conftest import synth_vrptw
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    inst, node, _ = synth_vrptw(31, 60, 5)
    final, keys, cost = O.local_search(inst, node)
    assert out[0] == out[1] == (len(keys), cost, final.tolist())


@pytest.mark.timeout(300)
def test_batched_replicated_search_under_gloo_matches_oracle():
    """SPEC §8.1 over two ranks: shard tables MIN-reduced through gloo, both ranks apply the oracle's batches."""
    from oracle import oracle as O
    from conftest import synth_vrptw
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker_batched, args=(2, port, out), nprocs=2, join=True)
    inst, node, _ = synth_vrptw(32, 50, 5)
    final, keys, sizes, cost = O.local_search_batched(inst, node)
    assert sizes.max() > 1
    assert out[0] == out[1] == (len(keys), cost, final.tolist())


class OracleEtaEngine:
    """Stand-in for Engine.eta_score on the CPU: the scalar evaluator."""

    def __init__(self, blob):
        self.blob = blob

    def eta_score(self, distance_m, weather, traffic, weekday, hour, age):
        from oracle import oracle as O
        return O.eta_eval(self.blob, O.eta_features(distance_m, weather, traffic, weekday, hour, age))


def _worker_eta(rank, world, port, out):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    import torch.distributed as dist
    from test_eta import synth_model, synth_pairs
    from routest_b200.dist import eta_score_distributed
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        _, blob = synth_model(seed=7, n_estimators=20, depth=4, n=1500)
        cols = synth_pairs(11, 10_001)   # odd: the two ranges differ in length
        got = eta_score_distributed(OracleEtaEngine(blob), cols["distance_m"], cols["weather"], cols["traffic"],
                                    cols["weekday"], cols["hour"], cols["age"])
        out[rank] = got.numpy().tobytes()
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_eta_pairs_sharded_over_two_ranks_equal_the_single_rank_result():
    """contiguous pair ranges per rank, model replicated; the gathered result is the
    single-rank result on every rank."""
    from oracle import oracle as O
    from test_eta import synth_model, synth_pairs
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker_eta, args=(2, port, out), nprocs=2, join=True)
    _, blob = synth_model(seed=7, n_estimators=20, depth=4, n=1500)
    cols = synth_pairs(11, 10_001)
    want = O.eta_eval(blob, O.eta_features(**cols)).astype(np.float32).tobytes()
    assert out[0] == out[1] == want
