"""Kernel micro-benchmark for the matrix build (development aid; bench.py is the judged harness)."""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from routest_b200 import Engine  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cases", default="1001x64,1001x128,5001x1,10001x1,20001x1")
    ap.add_argument("--iters", type=int, default=20)
    a = ap.parse_args()
    eng = Engine(0)
    torch.cuda.set_device(0)
    s = torch.cuda.current_stream().cuda_stream
    for case in a.cases.split(","):
        n, B = (int(x) for x in case.split("x"))
        g = torch.Generator(device="cuda").manual_seed(1)
        lat = 14.4043 + (14.6656 - 14.4043) * torch.rand((B, n), dtype=torch.float64, device="cuda", generator=g)
        lon = 120.9634 + (121.0891 - 120.9634) * torch.rand((B, n), dtype=torch.float64, device="cuda", generator=g)
        xyz = torch.empty((B, 3, n), dtype=torch.float64, device="cuda")
        pitch = eng.matrix_pitch(n)
        cells = torch.empty((B, n, pitch, 2), dtype=torch.int32, device="cuda")
        eng.unit_vectors_dev(lat.data_ptr(), lon.data_ptr(), n, B, xyz.data_ptr(), s)
        for _ in range(3):
            eng.matrix_cells_dev(xyz.data_ptr(), n, B, cells.data_ptr(), stream=s)
        torch.cuda.synchronize()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(a.iters + 1)]
        ev[0].record()
        for i in range(a.iters):
            eng.matrix_cells_dev(xyz.data_ptr(), n, B, cells.data_ptr(), stream=s)
            ev[i + 1].record()
        torch.cuda.synchronize()
        ms = sorted(ev[i].elapsed_time(ev[i + 1]) for i in range(a.iters))
        med = ms[len(ms) // 2]
        cells_n = B * n * n
        print(json.dumps({"n": n, "batch": B, "ms_med": round(med, 4), "ms_min": round(ms[0], 4),
                          "Tcells_s": round(cells_n / med / 1e9, 4), "GBs_algo": round(cells_n * 8 / med / 1e6, 1),
                          "GBs_written": round(B * n * pitch * 8 / med / 1e6, 1)}))
    eng.close()


if __name__ == "__main__":
    main()
