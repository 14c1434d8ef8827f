"""Where the end-to-end plan call spends its time (development aid)."""
import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import *  # noqa
from routest_b200 import Engine, ALL_FAMILIES

eng = Engine(0); B, stops = 64, 1000; n = stops + 1
pr = eng.problem(n, B)
raw = [synth_coords(2345 + b, stops, 10) for b in range(B)]
lat = np.stack([r[0] for r in raw]); lon = np.stack([r[1] for r in raw]); demand = np.stack([r[2] for r in raw]); cap = np.array([r[3] for r in raw], np.int32)
pr.set_coords(lat, lon); pr.build_matrix(DIST_SCALE, TS_CAR)
st = np.zeros((B, n, 4), np.int32)
for b in range(B):
    _, t = pr.get_matrix(b); st[b, :, :3] = synth_windows(2345 + b, t[0, 1:], t[1:, 0]); st[b, :, 3] = demand[b]
pr.set_stops(st, cap); trips, status = pr.greedy(demand.astype(float), cap.astype(float), use_tw=True); pr.set_routes(trips)
node = np.zeros((B, pr.pstride), np.int32); npos = np.zeros(B, np.int32)
for b in range(B): node[b], npos[b] = positions(trips[b], pr.pstride)
def pinned(a):
    t = torch.from_numpy(np.ascontiguousarray(a)).pin_memory(); return t, t.numpy()
keep = [pinned(x) for x in (lat, lon, st, cap, np.full(B, 2 ** 62, np.int64), node, npos, np.zeros(B, np.uint64))]
h = [k[1] for k in keep]
s = torch.cuda.current_stream().cuda_stream
keys_d = torch.empty(B, dtype=torch.int64, device="cuda")
def T(fn, reps=50):
    for _ in range(5): fn()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(reps): fn()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / reps * 1e3
print("plan (host buffers, sync inside)      ms", round(T(lambda: pr.plan(h[0], h[1], h[2], h[3], h[4], h[5], h[6], dist_scale=DIST_SCALE, time_scale=TS_CAR, out=h[7])), 4))
print("plan_dev + sync each call             ms", round(T(lambda: (pr.plan_dev(keys_d.data_ptr(), dist_scale=DIST_SCALE, time_scale=TS_CAR, stream=s), torch.cuda.synchronize())), 4))
print("plan_dev back to back                 ms", round(T(lambda: pr.plan_dev(keys_d.data_ptr(), dist_scale=DIST_SCALE, time_scale=TS_CAR, stream=s)), 4))
devs = [torch.empty_like(k[0], device="cuda") for k in keep[:7]]
print("7 H2D copies (torch, pinned) + sync   ms", round(T(lambda: ([d.copy_(k[0], non_blocking=True) for d, k in zip(devs, keep)], torch.cuda.synchronize())), 4))
big = torch.empty(sum(x.nbytes for x in h[:7]), dtype=torch.uint8).pin_memory(); bigd = torch.empty_like(big, device="cuda")
print("1 H2D copy of the same bytes + sync   ms", round(T(lambda: (bigd.copy_(big, non_blocking=True), torch.cuda.synchronize())), 4))
print("empty sync                            ms", round(T(lambda: torch.cuda.synchronize()), 4))
# two problems used in turn, inputs already on the device: does alternating cost anything by itself?
pr2 = eng.problem(n, B); pr2.set_coords(lat, lon); pr2.build_matrix(DIST_SCALE, TS_CAR); pr2.set_stops(st, cap); pr2.set_routes(trips)
keys2 = torch.empty(B, dtype=torch.int64, device="cuda")
flip = [0]
def alt():
    flip[0] ^= 1
    (pr2 if flip[0] else pr).plan_dev((keys2 if flip[0] else keys_d).data_ptr(), dist_scale=DIST_SCALE, time_scale=TS_CAR, stream=s)
print("plan_dev back to back, two problems   ms", round(T(alt, 100), 4))
outs = [pinned(np.zeros(B, np.uint64)) for _ in range(2)]
pipes = [p_.plan_pipeline_calls(h[0], h[1], h[2], h[3], h[4], h[5], h[6], o_[1], dist_scale=DIST_SCALE, time_scale=TS_CAR) for p_, o_ in ((pr, outs[0]), (pr2, outs[1]))]
def pipe(steps):
    pipes[0][0]()
    for k in range(1, steps):
        pipes[k % 2][0](); pipes[(k - 1) % 2][1]()
    pipes[(steps - 1) % 2][1]()
pipe(10); torch.cuda.synchronize(); t0 = time.perf_counter(); pipe(200); dt = (time.perf_counter() - t0) / 200 * 1e3
print("begin/end, two requests in flight     ms", round(dt, 4))
