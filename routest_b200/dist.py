"""Multi-GPU plumbing: one process per GPU (torchrun), torch.distributed for the only exchange the path has.

Where the path shards:
  * independent instances / ETA pairs / matrix rows  -> contiguous ranges per rank, no collective
    (`shard_range`, `eta_score_shard_dev`, `eta_score_distributed`);
  * local search on one large instance               -> every rank keeps a replica of the problem, scores its shard
    of the candidate space, and the ranks all-reduce ONE int64 key per instance with MIN; every rank then applies
    the same winning move (`DistributedSearch`).  Keys travel in order-preserving signed form (packed key XOR 2^63),
    so `ReduceOp.MIN` on int64 picks the lowest (delta, family, p, q) and INT64_MAX means "no move".
    With `batched=True` an iteration applies a whole set of independent moves (SPEC.md §8.1): every rank scores
    its shard into a table of best keys per ordered route pair, the ranks MIN-reduce that table (R^2 + R int64 per
    instance) — one all-reduce, or an all-gather into peer memory plus a local minimum — and every rank selects and
    applies the same batch.
  * greedy construction                               -> replicas only (sequential chain).
The reduce is 8 bytes per instance and latency bound.  The host does not pace the kernels one by one: everything
between two all-reduces (apply the winner, rebuild this rank's matrix rows and the route state, score the shard) is
captured ONCE into a CUDA graph and replayed, so an iteration costs the host one graph launch plus one collective
call, and the done flags are looked at only every `check_every` iterations.

Two transports for that reduce:
  exchange="nccl"  one `all_reduce(MIN)` per iteration through torch.distributed;
  exchange="peer"  no collective call in the loop: the replicas export a slot buffer through CUDA IPC once, and the
                   evaluation step MIN-reduces its keys straight into every peer's buffer with system-scope atomics
                   over NVLink (routest_b200/csrc/exchange.cu); torch.distributed only carries the handles at set-up.
                   The batched table travels the same way (every rank stores its shard table into every replica's
                   buffer, each replica takes the minimum locally).  Two iterations (the two slots) form one graph.
                   An NCCL all-reduce costs 18 us on two B200s and 29-40 us on eight whatever its size up to 0.5 MB;
                   the peer exchange about a third of that.
"""
from __future__ import annotations

import torch
import torch.distributed as dist

INT64_NONE = 2 ** 63 - 1
DONE_ABORTED = 3  # rt_problem_search_status: a replica gave up waiting for a peer (exchange.cu)


class PeerExchangeUnavailable(RuntimeError):
    """CUDA IPC export or mapping failed on some rank.  Raised on EVERY rank of the group together (the ranks agree on
    it through the handles they gather and one flag all-reduce), so callers can fall back to exchange="nccl" in step."""


class ReplicaAborted(RuntimeError):
    """A replica of a peer-exchange search stopped waiting for another one (dead or stalled peer): the replicas'
    routes can no longer be assumed identical, so the search has no result."""


def shard_range(total: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous, balanced split of `total` units: rank r gets [total*r/world, total*(r+1)/world)."""
    return total * rank // world, total * (rank + 1) // world


def allreduce_min_keys(keys: torch.Tensor, group=None) -> torch.Tensor:
    """In-place MIN all-reduce of signed-order move keys (int64, one per instance)."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(keys, op=dist.ReduceOp.MIN, group=group)
    return keys


class DistributedSearch:
    """Best-improvement local search over replicas.  `problem` needs the stream-ordered steps of the C ABI:
    eval_moves_dev(keys_ptr, families, shard, stream) and apply_moves_dev(keys_ptr, improving_only, stream), plus
    search_status() -> (done, moves, cost).  The engine's Problem provides them; the CPU tests substitute a stand-in
    to exercise this loop under gloo (no graphs there)."""

    def __init__(self, problem, batch: int, *, device=None, group=None, families: int = 7, check_every: int = 16,
                 exchange: str = "nccl", batched: bool = False, graph: bool = True):
        self.problem, self.batch, self.group = problem, batch, group
        self.families, self.check_every = families, check_every
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.keys = torch.full((batch,), INT64_NONE, dtype=torch.int64, device=device)
        if hasattr(problem, "set_shard"):  # the replica then keeps only its shard's rows of the route-order matrix fresh
            problem.set_shard(self.rank, self.world)
        if exchange not in ("nccl", "peer"):
            raise ValueError("exchange must be 'nccl' or 'peer'")
        self.exchange = exchange
        self.batched = batched
        # one spare entry: the peer exchange moves the table in 16-byte pairs
        self.table = (torch.full((problem.table_entries() + 1,), INT64_NONE, dtype=torch.int64, device=device)
                      if batched else None)
        self.use_graph = bool(graph) and self.keys.is_cuda
        self._graph = None       # everything between two all-reduces (nccl) / two whole iterations (peer)
        self._primed = False     # nccl: the first evaluation has been enqueued
        self._capture_refused = False
        if exchange == "peer":
            try:
                handle = problem.table_exchange_export(self.world) if batched else problem.exchange_export()
            except Exception:
                handle = None
            handles = [None] * self.world
            if self.world > 1:
                dist.all_gather_object(handles, handle, group=group)
            else:
                handles[0] = handle
            if any(h is None for h in handles):
                raise PeerExchangeUnavailable("CUDA IPC export failed on a rank")
            ok = True
            try:
                if batched:
                    problem.table_exchange_connect(handles, self.rank, self.world)
                else:
                    problem.exchange_connect(handles, self.rank, self.world)
            except Exception:
                ok = False
            if self.world > 1:   # also the barrier: nobody publishes into a buffer that is not mapped yet
                flag = torch.tensor([1 if ok else 0], device=self.keys.device)
                dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
                ok = bool(flag.item())
            if not ok:
                raise PeerExchangeUnavailable("mapping a peer's exchange buffer failed on a rank")

    def _stream(self) -> int:
        return torch.cuda.current_stream().cuda_stream if self.keys.is_cuda else 0

    # -- the pieces of an iteration, stream-ordered ---------------------------------------------------------------
    def _evaluate(self, s):
        if self.batched and self.exchange == "peer":
            self.problem.eval_table_xchg_dev(self.table.data_ptr(), self.families, s)
        elif self.batched:
            self.problem.eval_table_dev(self.table.data_ptr(), self.families, (self.rank, self.world), s)
        elif self.exchange == "peer":
            self.problem.eval_moves_xchg_dev(self.families, s)
        else:
            self.problem.eval_moves_dev(self.keys.data_ptr(), self.families, (self.rank, self.world), s)

    def _reduce(self):
        if self.exchange == "nccl":
            allreduce_min_keys(self.table if self.batched else self.keys, self.group)

    def _apply(self, s):
        if self.batched and self.exchange == "peer":
            self.problem.apply_batch_xchg_dev(self.table.data_ptr(), s)
        elif self.batched:
            self.problem.apply_batch_dev(self.table.data_ptr(), s)
        elif self.exchange == "peer":
            self.problem.apply_moves_xchg_dev(True, s)
        else:
            self.problem.apply_moves_dev(self.keys.data_ptr(), True, s)

    def step(self):
        """One iteration, kernel by kernel (no graph): evaluate the shard, combine, apply the winner everywhere."""
        s = self._stream()
        self._evaluate(s)
        self._reduce()
        self._apply(s)

    # -- graph replay ----------------------------------------------------------------------------------------------
    def _capture(self):
        """nccl: graph = apply(i) + refresh + evaluate(i+1), the collective stays an ordinary call between replays.
        peer: graph = two whole iterations (slot 0, slot 1), the exchange kernels are part of it."""
        g = torch.cuda.CUDAGraph()
        try:
            with torch.cuda.graph(g, capture_error_mode="thread_local"):
                s = torch.cuda.current_stream().cuda_stream
                if self.exchange == "peer":
                    for _ in range(2):
                        self._evaluate(s)
                        self._apply(s)
                else:
                    self._apply(s)
                    self._evaluate(s)
            self._graph = g
        except Exception:  # capture refused (driver, allocator, a stand-in problem): launch kernel by kernel instead
            self._graph = None
            self._capture_refused = True
            if self.exchange == "peer":
                self.use_graph = False
            torch.cuda.synchronize()

    def _run_chunk(self, iterations: int) -> int:
        """Enqueues about `iterations` iterations; returns how many were enqueued."""
        if not self.use_graph:
            for _ in range(iterations):
                self.step()
            return iterations
        s = self._stream()
        if self.exchange == "peer":
            by_hand = 0
            if self._graph is None:
                self.step()              # lazy set-up (side streams, buffers) happens outside the capture,
                self.step()              # and the capture starts on slot 0 again
                by_hand = 2
                self._capture()
                if self._graph is None:
                    for _ in range(max(iterations - by_hand, 0)):
                        self.step()
                    return max(iterations, by_hand)
            pairs = (max(iterations - by_hand, 0) + 1) // 2
            for _ in range(pairs):
                self._graph.replay()
            return by_hand + 2 * pairs
        if not self._primed:             # the evaluation whose keys the first all-reduce combines
            self._evaluate(s)
            self._primed = True
        by_hand = 0
        if self._graph is None and not self._capture_refused:
            self._reduce()               # one iteration by hand first, for the same reason
            self._apply(s)
            self._evaluate(s)
            by_hand = 1
            self._capture()
        for _ in range(max(iterations - by_hand, 0)):
            self._reduce()
            if self._graph is not None:
                self._graph.replay()
            else:                        # capture was refused: the same two steps, kernel by kernel
                self._apply(s)
                self._evaluate(s)
        return max(iterations, by_hand)

    def run(self, max_moves: int = 1 << 30):
        """Runs to a local optimum (or max_moves iterations).  Returns (moves, cost) per instance — identical on
        every rank.  Raises ReplicaAborted when a replica of a peer-exchange search stopped waiting for a peer."""
        issued = 0
        while issued < max_moves:
            todo = min(self.check_every, max_moves - issued)
            issued += self._run_chunk(todo)
            done, moves, cost = self.problem.search_status()
            if any(int(d) == DONE_ABORTED for d in done):
                raise ReplicaAborted(f"rank {self.rank}: a replica stopped waiting for a peer after {issued} iterations")
            if all(int(d) != 0 for d in done):
                break
        done, moves, cost = self.problem.search_status()
        if any(int(d) == DONE_ABORTED for d in done):
            raise ReplicaAborted(f"rank {self.rank}: a replica stopped waiting for a peer")
        return moves, cost


# ---- ETA pairs: contiguous pair ranges per rank, the model replicated, no collective on the
# data path; the gather of the outputs is a concat.

_ETA_ITEM = (8, 1, 1, 1, 1, 4)  # bytes per pair of distance_m, weather, traffic, weekday, hour, driver_age


def eta_score_shard_dev(engine, col_ptrs, n_pairs: int, out_ptr: int, rank: int, world: int, stream: int = 0):
    """Scores this rank's contiguous range of `n_pairs` stop pairs (rt_eta_score_dev on offset device pointers).
    col_ptrs: device pointers of the six full-length input columns in the order of include/routest.h; out_ptr: the
    full-length float32 output.  Returns the range [lo, hi) that was written."""
    lo, hi = shard_range(n_pairs, rank, world)
    if hi > lo:
        ptrs = [int(p) + lo * w for p, w in zip(col_ptrs, _ETA_ITEM)]
        engine.eta_score_dev(*ptrs, hi - lo, int(out_ptr) + 4 * lo, stream)
    return lo, hi


def eta_score_distributed(engine, distance_m, weather, traffic, weekday, hour, age, group=None) -> torch.Tensor:
    """Host arrays in, minutes for EVERY pair out on every rank (float32 tensor): each rank scores its own range
    with engine.eta_score and the ranges are gathered (ml.py:35-53, batched and sharded).  With one rank (or no
    process group) it is a plain engine.eta_score call."""
    import numpy as np
    n = len(distance_m)
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    lo, hi = shard_range(n, rank, world)
    cols = [np.asarray(c)[lo:hi] for c in (distance_m, weather, traffic, weekday, hour, age)]
    mine = np.asarray(engine.eta_score(*cols), dtype=np.float32) if hi > lo else np.zeros(0, np.float32)
    if world == 1:
        return torch.from_numpy(mine.copy())
    width = max(shard_range(n, r, world)[1] - shard_range(n, r, world)[0] for r in range(world))
    backend = dist.get_backend(group)
    dev = torch.device("cuda", torch.cuda.current_device()) if backend == "nccl" else torch.device("cpu")
    pad = torch.zeros(width, dtype=torch.float32, device=dev)
    pad[:hi - lo] = torch.from_numpy(mine.copy()).to(dev)
    parts = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(parts, pad, group=group)
    out = torch.cat([parts[r][:shard_range(n, r, world)[1] - shard_range(n, r, world)[0]] for r in range(world)])
    return out.cpu()
