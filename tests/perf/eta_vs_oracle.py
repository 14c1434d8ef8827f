"""ETA scorer micro-benchmark (BASELINE config 3: 1M synthetic stop pairs; development aid)."""
import json, os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from routest_b200 import Engine
from test_eta import synth_model, synth_pairs
from oracle import oracle as O

eng = Engine(0)
n = 1_000_000
for trees, depth in ((120, 6), (500, 6)):
    _, blob = synth_model(seed=3, n_estimators=trees, depth=depth)
    eng.eta_load_model(blob)
    nt, nn = eng.eta_model_info()
    cols = synth_pairs(3456, n)
    dev = {k: torch.tensor(v, device="cuda") for k, v in cols.items()}
    out = torch.empty(n, dtype=torch.float32, device="cuda")
    s = torch.cuda.current_stream().cuda_stream
    run = lambda: eng.eta_score_dev(dev["distance_m"].data_ptr(), dev["weather"].data_ptr(), dev["traffic"].data_ptr(),
                                    dev["weekday"].data_ptr(), dev["hour"].data_ptr(), dev["age"].data_ptr(), n, out.data_ptr(), s)
    for _ in range(3): run()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(10): run()
    b.record(); torch.cuda.synchronize()
    ms = a.elapsed_time(b) / 10
    t0 = time.perf_counter(); host = eng.eta_score(**cols); e2e = time.perf_counter() - t0
    rows = O.eta_features(**{k: v[:200000] for k, v in cols.items()})
    t0 = time.perf_counter(); ref = O.eta_eval(blob, rows, threads=O.num_threads()); cpu = time.perf_counter() - t0
    print(json.dumps({"trees": nt, "nodes": nn, "pairs": n, "ms_device": round(ms, 4), "Gpairs_s_device": round(n / ms / 1e6, 3),
                      "ms_host_api": round(e2e * 1e3, 2), "Mpairs_s_host_api": round(n / e2e / 1e6, 1),
                      "cpu_Mpairs_s": round(200000 / cpu / 1e6, 2), "cpu_threads": O.num_threads(),
                      "bit_exact_sample": bool(np.array_equal(host[:200000], ref))}))
