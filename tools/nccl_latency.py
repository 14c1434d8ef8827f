"""NCCL all-reduce (MIN, int64) latency for the sizes this path exchanges, device time per call, max over ranks
(development aid, torchrun).  Environment variables such as NCCL_ALGO / NCCL_PROTO select what is measured."""
import json
import os

import torch
import torch.distributed as dist


def main():
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    out = {"world": world, "NCCL_ALGO": os.environ.get("NCCL_ALGO"), "NCCL_PROTO": os.environ.get("NCCL_PROTO"),
           "NCCL_NVLS_ENABLE": os.environ.get("NCCL_NVLS_ENABLE")}
    for name, cnt in (("8B", 1), ("20KB", 2550), ("551KB", 68906)):
        t = torch.zeros(cnt, dtype=torch.int64, device="cuda")
        for _ in range(20):
            dist.all_reduce(t, op=dist.ReduceOp.MIN)
        torch.cuda.synchronize()
        dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(300):
            dist.all_reduce(t, op=dist.ReduceOp.MIN)
        e1.record()
        torch.cuda.synchronize()
        us = torch.tensor([e0.elapsed_time(e1) / 300 * 1e3], device="cuda")
        dist.all_reduce(us, op=dist.ReduceOp.MAX)
        out[f"allreduce_{name}_us"] = round(float(us.item()), 2)
    if rank == 0:
        print(json.dumps(out), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
