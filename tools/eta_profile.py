"""A few scoring calls of 1M pairs on the bench's 500-tree depth-6 ensemble (for ncu; development aid)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from routest_b200 import Engine  # noqa: E402

eng = Engine(0)
eng.eta_load_model(bench.eta_model_blob(500, 6))
m = 1_000_000
cols = [torch.tensor(c, device="cuda") for c in bench.eta_columns(m)]
res = torch.empty(m, dtype=torch.float32, device="cuda")
for _ in range(3):
    eng.eta_score_dev(*[c.data_ptr() for c in cols], m, res.data_ptr(), torch.cuda.current_stream().cuda_stream)
torch.cuda.synchronize()
print(float(res.sum()))
