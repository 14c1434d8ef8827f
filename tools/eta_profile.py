"""One scoring call of 1M pairs on a synthetic 200-tree depth-6 ensemble (for ncu; development aid)."""
import os, struct, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from routest_b200 import Engine

eng = Engine(0)
rng = np.random.default_rng(3456)
trees, depth = 200, 6
nodes, offs = bytearray(), [0]
for _ in range(trees):
    cnt = 2 ** (depth + 1) - 1
    for k in range(cnt):
        if k < 2 ** depth - 1:
            f = int(rng.integers(0, 12))
            thr = float(rng.uniform(0, 40) if f == 10 else rng.uniform(20, 40) if f == 11 else rng.uniform(0, 23) if f == 9
                        else rng.uniform(0, 6) if f == 8 else 0.5)
            nodes += struct.pack("<fHBB", thr, 2 * k + 1, f, 0)
        else:
            nodes += struct.pack("<fHBB", float(rng.normal(0, 0.2)), 0, 0, 1)
    offs.append(offs[-1] + cnt)
eng.eta_load_model(b"RTE1" + struct.pack("<IIf", trees, 12, 25.0) + struct.pack(f"<{len(offs)}I", *offs) + bytes(nodes))
m = 1_000_000
c = dict(d=torch.tensor(rng.uniform(0, 45000, m), device="cuda"), w=torch.tensor(rng.integers(0, 4, m).astype(np.uint8), device="cuda"),
         t=torch.tensor(rng.integers(0, 4, m).astype(np.uint8), device="cuda"), wd=torch.tensor(rng.integers(0, 7, m).astype(np.uint8), device="cuda"),
         h=torch.tensor(rng.integers(0, 24, m).astype(np.uint8), device="cuda"), a=torch.tensor(rng.uniform(20, 40, m).astype(np.float32), device="cuda"))
res = torch.empty(m, dtype=torch.float32, device="cuda")
for _ in range(3):
    eng.eta_score_dev(c["d"].data_ptr(), c["w"].data_ptr(), c["t"].data_ptr(), c["wd"].data_ptr(), c["h"].data_ptr(), c["a"].data_ptr(), m,
                      res.data_ptr(), torch.cuda.current_stream().cuda_stream)
torch.cuda.synchronize()
print(float(res.sum()))
