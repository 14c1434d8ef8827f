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
