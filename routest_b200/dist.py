"""Multi-GPU plumbing: one process per GPU (torchrun), torch.distributed for the only exchange the path has.

Where the path shards:
  * independent instances / ETA pairs / matrix rows  -> contiguous ranges per rank, no collective
    (`shard_range`);
  * local search on one large instance               -> every rank keeps a replica of the problem, scores its shard
    of the candidate space, and the ranks all-reduce ONE int64 key per instance with MIN; every rank then applies
    the same winning move (`DistributedSearch`).  Keys travel in order-preserving signed form (packed key XOR 2^63),
    so `ReduceOp.MIN` on int64 picks the lowest (delta, family, p, q) and INT64_MAX means "no move".
    With `batched=True` an iteration applies a whole set of independent moves (SPEC.md §8.1): every rank scores
    its shard into a table of best keys per ordered route pair, the ranks MIN-reduce that table (R^2 + R int64 per
    instance) in one all-reduce, and every rank selects and applies the same batch.
  * greedy construction                               -> replicas only (sequential chain).
The reduce is 8 bytes per instance and latency bound; iterations are enqueued back to back on one stream and the
host looks at the done flags only every `check_every` iterations.

Two transports for that reduce:
  exchange="nccl"  one `all_reduce(MIN)` per iteration through torch.distributed;
  exchange="peer"  no collective call in the loop: the replicas export a slot buffer through CUDA IPC once, and the
                   evaluation step MIN-reduces its keys straight into every peer's buffer with system-scope atomics
                   over NVLink (routest_b200/csrc/exchange.cu); torch.distributed only carries the handles at set-up.
"""
from __future__ import annotations

import torch
import torch.distributed as dist

INT64_NONE = 2 ** 63 - 1


def shard_range(total: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous, balanced split of `total` units: rank r gets [total*r/world, total*(r+1)/world)."""
    return total * rank // world, total * (rank + 1) // world


def allreduce_min_keys(keys: torch.Tensor, group=None) -> torch.Tensor:
    """In-place MIN all-reduce of signed-order move keys (int64, one per instance)."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(keys, op=dist.ReduceOp.MIN, group=group)
    return keys


class DistributedSearch:
    """Best-improvement local search over replicas.  `problem` needs the two stream-ordered steps of the C ABI:
    eval_moves_dev(keys_ptr, families, shard, stream) and apply_moves_dev(keys_ptr, improving_only, stream), plus
    search_status() -> (done, moves, cost).  The engine's Problem provides them; the CPU tests substitute a stand-in
    to exercise this loop under gloo."""

    def __init__(self, problem, batch: int, *, device=None, group=None, families: int = 7, check_every: int = 16,
                 exchange: str = "nccl", batched: bool = False):
        self.problem, self.batch, self.group = problem, batch, group
        self.families, self.check_every = families, check_every
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.keys = torch.full((batch,), INT64_NONE, dtype=torch.int64, device=device)
        if hasattr(problem, "set_shard"):  # the replica then keeps only its shard's rows of the route-order matrix fresh
            problem.set_shard(self.rank, self.world)
        if exchange not in ("nccl", "peer"):
            raise ValueError("exchange must be 'nccl' or 'peer'")
        if batched and exchange != "nccl":
            raise ValueError("batched iterations exchange their table through the collective (exchange='nccl')")
        self.exchange = exchange
        self.batched = batched
        self.table = (torch.empty(problem.table_entries(), dtype=torch.int64, device=device) if batched else None)
        if exchange == "peer":
            handle = problem.exchange_export()
            handles = [None] * self.world
            if self.world > 1:
                dist.all_gather_object(handles, handle, group=group)
            else:
                handles[0] = handle
            problem.exchange_connect(handles, self.rank, self.world)
            if self.world > 1:
                dist.barrier(group=group)  # nobody publishes into a buffer that is not mapped yet

    def _stream(self) -> int:
        return torch.cuda.current_stream().cuda_stream if self.keys.is_cuda else 0

    def step(self):
        s = self._stream()
        if self.batched:
            self.problem.eval_table_dev(self.table.data_ptr(), self.families, (self.rank, self.world), s)
            allreduce_min_keys(self.table, self.group)
            self.problem.apply_batch_dev(self.table.data_ptr(), s)
            return
        if self.exchange == "peer":
            self.problem.eval_moves_xchg_dev(self.families, s)
            self.problem.apply_moves_xchg_dev(True, s)
            return
        self.problem.eval_moves_dev(self.keys.data_ptr(), self.families, (self.rank, self.world), s)
        allreduce_min_keys(self.keys, self.group)
        self.problem.apply_moves_dev(self.keys.data_ptr(), True, s)

    def run(self, max_moves: int = 1 << 30):
        """Runs to a local optimum (or max_moves iterations).  Returns (moves, cost) per instance — identical on
        every rank."""
        issued = 0
        while issued < max_moves:
            todo = min(self.check_every, max_moves - issued)
            for _ in range(todo):
                self.step()
            issued += todo
            done, moves, cost = self.problem.search_status()
            if all(int(d) != 0 for d in done):
                break
        done, moves, cost = self.problem.search_status()
        return moves, cost
