"""Where an iteration of the replicated search goes on N GPUs (development aid, torchrun): the collective alone, the
graph-replayed segments alone, and both together — device time per iteration, max over ranks.

  python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 --master-port 29561 tools/coll_latency.py
"""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from routest_b200 import ALL_FAMILIES, Engine  # noqa: E402


def main():
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    eng = Engine(local)

    def timed(fn, reps=200):
        for _ in range(10):
            fn()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / reps * 1e3], device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    out = {"world": world}
    for name, cnt in (("allreduce_8B_us", 1), ("allreduce_20KB_us", 2550), ("allreduce_551KB_us", 68906)):
        t = torch.zeros(cnt, dtype=torch.int64, device="cuda")
        out[name] = timed(lambda: dist.all_reduce(t, op=dist.ReduceOp.MIN)) if world > 1 else 0.0
    # batched iteration pieces on the 5,000-stop instance
    pr, w = bench.build_problem(eng, bench.SEED_5K, 5000, 50)
    pr.set_shard(rank, world)
    table = torch.empty(pr.table_entries(), dtype=torch.int64, device="cuda")
    s = torch.cuda.current_stream().cuda_stream
    out["table_entries"] = table.numel()
    out["eval_table_us"] = timed(lambda: pr.eval_table_dev(table.data_ptr(), ALL_FAMILIES, (rank, world), s))
    # apply on a table with no improving entry is a no-op batch; time the refresh it ends with instead
    out["refresh_us"] = timed(lambda: pr.refresh_dev(3, s))
    replay, graphed = bench.capture(torch, lambda st: (pr.refresh_dev(3, st),
                                                       pr.eval_table_dev(table.data_ptr(), ALL_FAMILIES, (rank, world), st)))
    out["graph_refresh_eval_us"] = timed(replay)
    if world > 1:
        out["graph_plus_allreduce_us"] = timed(lambda: (replay(), dist.all_reduce(table, op=dist.ReduceOp.MIN)))
    if rank == 0:
        print(json.dumps(out))
    pr.close()
    eng.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
