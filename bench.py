#!/usr/bin/env python
"""bench.py — one JSON line for the routing hot path on N B200s (contract in the task statement).

A step = one pass of the hot path over one batch of synthetic input: for each of `--batch` independent
1,000-stop VRPTW instances (BASELINE.json configs[1]) build the 1,001 x 1,001 distance/time matrix from the
coordinates (in route order: one row per position of the current solution), derive the route state of the current solution and score the full 2-opt/relocate/swap neighbourhood
(delta + capacity + time-window feasibility per candidate), ending in one best-move key per instance.

  value        matrix cells/s over the whole step, inputs resident in HBM, CUDA events on the launching stream
  e2e          the same metric through rt_problem_plan with pinned HOST buffers (H2D of coordinates, stops and the
               current routes, D2H of the best keys, inside the timed region)
  roofline     the step's dominant kernel, timed alone in the same process
  cpu_baseline the CPU oracle (a port: the reference has no code for these steps) on this box's cores
At N > 1 every rank owns its own batch (independent instances, no data-path collective): weak scaling.

`--impl reference` times the CPU oracle instead (rank 0 only) and prints the same line with "impl": "reference".
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

BOX = (14.4043, 14.6656, 120.9634, 121.0891)
DEPOT = (14.5836, 121.0409)
TS_CAR = 3.6 / 40.0
DIST_SCALE = 10.0
METRIC = "matrix cells/s and VRPTW move-evals/s"
BYTES_PER_EVAL = {0: 16, 1: 24, 2: 32}  # matrix reads per 2-opt / relocate / swap candidate


# ----------------------------------------------------------------------------------------------- workload

def synth_coords(seed, stops, vehicles):
    rng = np.random.default_rng(seed)
    lat = np.concatenate([[DEPOT[0]], rng.uniform(BOX[0], BOX[1], stops)])
    lon = np.concatenate([[DEPOT[1]], rng.uniform(BOX[2], BOX[3], stops)])
    demand = np.concatenate([[0], rng.integers(1, 10, stops)]).astype(np.int64)
    cap = int(np.ceil(1.05 * demand.sum() / vehicles))
    return lat, lon, demand, cap


def synth_windows(seed, t_out, t_back, horizon=43200, service=300):
    n_stops = len(t_out)
    rng = np.random.default_rng(seed + 7919)
    centre = rng.integers(0, horizon + 1, n_stops)
    half = rng.integers(900, 3601, n_stops)
    centre = np.clip(centre, t_out, horizon - service - np.asarray(t_back))
    tw = np.zeros((n_stops + 1, 3), np.int32)
    tw[0] = (0, horizon, 0)
    tw[1:, 0] = np.clip(centre - half, 0, horizon)
    tw[1:, 1] = np.clip(centre + half, 0, horizon)
    tw[1:, 2] = service
    return tw


def positions(trips, pstride):
    node = np.zeros(pstride, np.int32)
    p = 1
    for t in trips:
        node[p:p + len(t)] = t
        p += len(t) + 1
    return node, p


def count_candidates(trips, n):
    """Valid candidates per family (SPEC.md §7) — the unit of move-evals/s."""
    ncust = n - 1
    P = n + len(trips)
    two_opt = sum(len(t) * (len(t) - 1) // 2 for t in trips)
    return {0: two_opt, 1: ncust * (P - 3), 2: ncust * (ncust - 1) // 2}


# ----------------------------------------------------------------------------------------------- clocks

class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu, self.rows, self.proc = gpu_index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.gpu}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), [x.strip() for x in line.split(",")]))

    def stop(self, t0, t1):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        rows = [r for t, r in self.rows if t0 <= t <= t1 + 0.15] or [r for _, r in self.rows[-3:]]
        sm = [float(r[1]) for r in rows if r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in rows if r[2].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in rows for i in range(4) if len(r) > 4 + i and r[4 + i].lower() == "active"})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(rows)}


# ----------------------------------------------------------------------------------------------- CPU side

def host_threads() -> int:
    """Threads for the CPU legs: every core this process may run on.  Not OMP_NUM_THREADS — torchrun sets that to 1
    for its workers, which would time a single-threaded baseline at N > 1."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def oracle_instances(seeds, stops, vehicles):
    """Instances + greedy start on the CPU oracle (used by the cpu_baseline leg and by --impl reference only)."""
    from oracle import oracle as O
    out = []
    for seed in seeds:
        lat, lon, demand, cap = synth_coords(seed, stops, vehicles)
        n = stops + 1
        xyz = O.unit_vectors(lat, lon)
        m = O.matrix_rows(xyz, 0, n, DIST_SCALE, TS_CAR, threads=host_threads())
        tw = synth_windows(seed, m["time"][0, 1:], m["time"][1:, 0])
        st = np.zeros((n, 4), np.int32)
        st[:, :3] = tw
        st[:, 3] = demand
        rc, trips = O.greedy_trips(m["dist"], demand.astype(float), float(cap), 9e12, T=m["time"], tw=tw)
        assert rc == 0
        out.append(dict(lat=lat, lon=lon, stops=st, cap=cap, trips=trips, node=O.build_positions(trips, n)))
    return out


def oracle_step(insts):
    """One pass of the hot path on the CPU oracle with all host threads; returns (cells, evals)."""
    from oracle import oracle as O
    th = host_threads()
    cells = evals = 0
    for it in insts:
        n = it["lat"].size
        m = O.matrix_rows(O.unit_vectors(it["lat"], it["lon"]), 0, n, DIST_SCALE, TS_CAR, threads=th)
        inst = O.Instance(m["dist"], m["time"], it["stops"], it["cap"])
        O.best_move(inst, it["node"], threads=th)
        cells += n * n
        evals += sum(count_candidates(it["trips"], n).values())
    return cells, evals


def cpu_baseline(stops, vehicles, budget_s=8.0, max_inst=4):
    from oracle import oracle as O
    insts = oracle_instances([2345 + i for i in range(max_inst)], stops, vehicles)
    oracle_step(insts[:1])  # warm-up
    t0 = time.perf_counter()
    cells = evals = reps = 0
    while time.perf_counter() - t0 < budget_s:
        c, e = oracle_step(insts)
        cells, evals, reps = cells + c, evals + e, reps + 1
    dt = time.perf_counter() - t0
    return {"value": cells / dt, "unit": "cells/s", "move_evals_per_s": evals / dt, "cores": host_threads(),
            "kind": "port",
            "sample": f"{reps} passes over {max_inst} of the workload's instances ({stops} stops each): oracle matrix "
                      f"build + full-neighbourhood best move, {host_threads()} OpenMP threads, {dt:.1f} s"}


def run_reference(args, rank):
    if rank != 0:
        return
    from oracle import oracle as O
    per_step = 2
    insts = oracle_instances([2345 + i for i in range(per_step)], args.stops, args.vehicles)
    for _ in range(args.warmup):
        oracle_step(insts)
    t0 = time.perf_counter()
    cells = evals = 0
    for _ in range(args.steps):
        c, e = oracle_step(insts)
        cells, evals = cells + c, evals + e
    dt = time.perf_counter() - t0
    val = cells / dt
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": "cells/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64/f32->i32", "data": "synthetic",
        "config": {"workload": f"{per_step} x ({args.stops}-stop matrix + full 2-opt/relocate/swap neighbourhood) per "
                               "step on the host CPU (bounded sample of the engine arm's batch)",
                   "stops": args.stops, "instances_per_step": per_step},
        "move_evals_per_s": evals / dt,
        "cpu_baseline": {"value": val, "unit": "cells/s", "cores": host_threads(), "kind": "port",
                         "sample": f"{args.steps} steps x {per_step} instances, CPU oracle (the reference itself has no "
                                   "code for matrix build or move evaluation: SURVEY.md §0)"},
        "e2e": {"value": val, "unit": "cells/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


# ----------------------------------------------------------------------------------------------- engine arm

def run_engine(args, rank, world):
    import torch
    import torch.distributed as dist

    from routest_b200 import ALL_FAMILIES, Engine

    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    eng = Engine(local)
    B, stops, n = args.batch, args.stops, args.stops + 1
    pr = eng.problem(n, B)

    # ---- set-up (untimed): this rank's instances, their windows and a feasible start from the engine's own greedy
    seeds = [2345 + rank * B + b for b in range(B)]
    raw = [synth_coords(s, stops, args.vehicles) for s in seeds]
    lat = np.stack([r[0] for r in raw])
    lon = np.stack([r[1] for r in raw])
    demand = np.stack([r[2] for r in raw])
    cap = np.array([r[3] for r in raw], np.int32)
    pr.set_coords(lat, lon)
    pr.build_matrix(DIST_SCALE, TS_CAR)
    st = np.zeros((B, n, 4), np.int32)
    for b in range(B):
        _, t = pr.get_matrix(b)
        st[b, :, :3] = synth_windows(seeds[b], t[0, 1:], t[1:, 0])
        st[b, :, 3] = demand[b]
    pr.set_stops(st, cap)
    trips, status = pr.greedy(demand.astype(np.float64), cap.astype(np.float64), use_tw=True)
    assert not status.any(), "greedy failed on the synthetic workload"
    pr.set_routes(trips)
    cand = [count_candidates(t, n) for t in trips]
    evals_step = sum(sum(c.values()) for c in cand)
    eval_bytes = sum(c[f] * BYTES_PER_EVAL[f] for c in cand for f in c)
    cells_step = B * n * n                       # the metric counts the n x n matrix of the instance ...
    pmax = max(n + len(t) for t in trips)
    cells_built = B * pmax * ((pmax + 15) // 16 * 16)   # ... the kernel writes it in route order with one depot row per route

    stream = torch.cuda.current_stream().cuda_stream
    keys_d = torch.empty(B, dtype=torch.int64, device="cuda")

    def step():
        # unit vectors -> matrix (route order) -> route state -> full neighbourhood -> best keys, all on `stream`
        pr.plan_dev(keys_d.data_ptr(), dist_scale=DIST_SCALE, time_scale=TS_CAR, families=ALL_FAMILIES, stream=stream)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    # ---- device-resident timing
    for _ in range(args.warmup):
        step()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.25)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t_wall0 = time.perf_counter()
    ev0.record()
    for _ in range(args.steps):
        step()
    ev1.record()
    barrier()
    t_wall1 = time.perf_counter()
    ms = ev0.elapsed_time(ev1)
    if world > 1:
        t = torch.tensor([ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    keys_ref = keys_d.clone()

    # ---- per-kernel timing (same process, alone): which one dominates, and its roofline
    def timed(fn, reps):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(reps):
            fn()
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) / reps

    reps = max(args.steps, 10)
    ms_matrix = timed(lambda: pr.refresh_dev(1, stream), reps)   # position vectors + matrix kernel (route order)
    ms_refresh = timed(lambda: pr.refresh_dev(2, stream), reps)  # gather + walk
    ms_eval = timed(lambda: pr.eval_moves_dev(keys_d.data_ptr(), ALL_FAMILIES, (0, 1), stream), reps)
    clocks = sampler.stop(t_wall0, time.perf_counter()) if rank == 0 else None

    # ---- end to end through the C ABI with pinned host buffers
    def pinned(a):
        t = torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
        return t, t.numpy()

    node = np.zeros((B, pr.pstride), np.int32)
    npos = np.zeros(B, np.int32)
    for b in range(B):
        node[b], npos[b] = positions(trips[b], pr.pstride)
    keep = [pinned(x) for x in (lat, lon, st, cap, np.full(B, 2 ** 62, np.int64), node, npos, np.zeros(B, np.uint64))]
    h = [k[1] for k in keep]
    h2d = sum(x.nbytes for x in h[:7])
    d2h = h[7].nbytes

    # the call a service makes per request batch: host buffers (re-filled in place by the caller) in, best keys out
    e2e_step = pr.plan_call(h[0], h[1], h[2], h[3], h[4], h[5], h[6], h[7], dist_scale=DIST_SCALE, time_scale=TS_CAR)

    for _ in range(args.warmup):
        e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        e2e_step()
    barrier()
    e2e_s = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([e2e_s], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    same = bool(np.array_equal(h[7].view(np.int64) ^ np.int64(-2 ** 63), keys_ref.cpu().numpy()))

    # ---- the same requests with two in flight (rt_problem_plan_begin / _end on two problems used in turn): every step
    # still uploads its inputs from pinned host memory and reads its keys back, but the upload of step k+1 runs under
    # the kernels of step k.  Reported next to the synchronous figure, not instead of it.
    pr2 = eng.problem(n, B)
    out2 = pinned(np.zeros(B, np.uint64))
    pipes = [p_.plan_pipeline_calls(h[0], h[1], h[2], h[3], h[4], h[5], h[6], o_, dist_scale=DIST_SCALE, time_scale=TS_CAR)
             for p_, o_ in ((pr, h[7]), (pr2, out2[1]))]

    def pipelined(steps):
        pipes[0][0]()
        for k in range(1, steps):
            pipes[k % 2][0]()
            pipes[(k - 1) % 2][1]()
        return pipes[(steps - 1) % 2][1]()

    pipelined(max(args.warmup, 2))
    barrier()
    t0 = time.perf_counter()
    last = pipelined(args.steps)
    barrier()
    pipe_s = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([pipe_s], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        pipe_s = float(t.item())
    same_pipe = bool(np.array_equal(last.view(np.int64) ^ np.int64(-2 ** 63), keys_ref.cpu().numpy()))
    pr2.close()

    extra = secondary_figures(eng, torch, stream) if (rank == 0 and not args.no_extra) else None

    if rank == 0:
        peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
        if os.path.exists(peaks_path):
            peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        else:
            peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
        traffic = None
        tp = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tp):
            traffic = json.load(open(tp))
        kern = {
            "matrix_kernel": {"ms": ms_matrix, "algorithmic_bytes": cells_step * 8,
                              "GBs": cells_step * 8 / ms_matrix / 1e6, "bytes_written": cells_built * 8,
                              "GBs_written": cells_built * 8 / ms_matrix / 1e6,
                              "note": "position_vectors_kernel + matrix_kernel<0> on the route-ordered stops"},
            "state_tables": {"ms": ms_refresh},
            "eval_moves(inter+intra)": {"ms": ms_eval, "algorithmic_bytes": eval_bytes,
                                        "GBs": eval_bytes / ms_eval / 1e6, "move_evals_per_s": evals_step / ms_eval * 1e3},
        }
        dom = "matrix_kernel" if ms_matrix >= ms_eval else "eval_moves(inter+intra)"
        ach = kern[dom]["GBs"]
        roof = {"kernel": dom, "bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                "peak_source": peak_src,
                "traffic": (traffic or {}).get(dom),
                "note": "algorithmic bytes per SURVEY.md §8(d): 8 B per matrix cell; 16/24/32 B of matrix reads per "
                        "2-opt/relocate/swap candidate (tile reuse makes real DRAM traffic much smaller than that)",
                "matrix_kernel": {"achieved": kern["matrix_kernel"]["GBs"], "frac": kern["matrix_kernel"]["GBs"] / peak}}
        # Neither kernel of this step is purely HBM-bound at 1,040-wide instances: both sit near the issue limit.  The
        # instruction counts are fixed by the workload (ncu, profiles/traffic.json); the rates use the live durations.
        winst = (traffic or {}).get("warp_instructions") or {}
        if winst and clocks.get("sm_mhz"):
            peak_issue = 148 * 4 * float(clocks["sm_mhz"]) * 1e6            # one warp instruction per SMSP per clock
            roof["issue"] = {k: {"warp_inst_per_launch": winst[k], "achieved_Ginst_s": winst[k] / kern[k]["ms"] / 1e6,
                                 "frac_of_issue_peak": winst[k] / (kern[k]["ms"] / 1e3) / peak_issue}
                             for k in ("matrix_kernel", "eval_moves(inter+intra)") if k in winst}
            roof["issue"]["peak_Ginst_s"] = peak_issue / 1e9
        total_cells = cells_step * world
        line = {
            "metric": METRIC, "value": total_cells * args.steps / (ms / 1e3), "unit": "cells/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64/f32->i32",
            "data": "synthetic",
            "config": {"workload": f"{B} x ({stops}-stop distance/time matrix + route state + full 2-opt/relocate/swap "
                                   "neighbourhood with capacity and time-window feasibility) per GPU per step",
                       "stops": stops, "instances_per_gpu": B, "routes_per_instance_mean": float(np.mean([len(t) for t in trips])),
                       "cells_per_step_per_gpu": cells_step, "move_evals_per_step_per_gpu": evals_step,
                       "l2": f"each step writes and re-reads {cells_step * 8 / 1e6:.0f} MB of matrix per GPU (> 126 MB L2); no flush needed",
                       "parallelism": f"{world} x independent batches (no collective)"},
            "move_evals_per_s": evals_step * world * args.steps / (ms / 1e3),
            "kernels": kern, "roofline": roof,
            "e2e": {"value": total_cells * args.steps / e2e_s, "unit": "cells/s", "h2d_bytes_per_step": int(h2d),
                    "d2h_bytes_per_step": int(d2h), "ms_per_step": 1e3 * e2e_s / args.steps,
                    "move_evals_per_s": evals_step * world * args.steps / e2e_s,
                    "call": "rt_problem_plan via Problem.plan_call (pinned host buffers in, best keys out; matrices stay in HBM)",
                    "keys_equal_device_path": same,
                    "two_requests_in_flight": {"value": total_cells * args.steps / pipe_s, "unit": "cells/s",
                                               "ms_per_step": 1e3 * pipe_s / args.steps,
                                               "call": "rt_problem_plan_begin / rt_problem_plan_end on two problems used in "
                                                       "turn: the upload of one request runs under the kernels of the other",
                                               "keys_equal_device_path": same_pipe}},
            "gpu_launches": 6 * args.steps,  # kernels per step, counted in profiles/r1_launches_bench_steps2.csv
            "clocks": clocks,
        }
        if extra:
            line["extra"] = extra
        if world == 1 and not args.no_cpu:
            line["cpu_baseline"] = cpu_baseline(stops, args.vehicles)
        print(json.dumps(line))
    pr.close()
    eng.close()
    if world > 1:
        dist.destroy_process_group()


def secondary_figures(eng, torch, stream):
    """Quick figures for the other BASELINE configs (not the headline): the matrix kernel at 10,001 nodes, the host
    matrix API, ETA scoring of 1M pairs on a synthetic 200-tree ensemble, and a full 5,000-stop solve."""
    import struct

    out = {}

    def timed(fn, reps=10):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(reps):
            fn()
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) / reps

    # config 4, matrix only: 10,001 nodes device-resident
    n = 10001
    lat, lon, _, _ = synth_coords(4567, n - 1, 100)
    lat_d, lon_d = torch.tensor(lat, device="cuda"), torch.tensor(lon, device="cuda")
    xyz = torch.empty((3, n), dtype=torch.float64, device="cuda")
    pitch = eng.matrix_pitch(n)
    cells = torch.empty((n, pitch, 2), dtype=torch.int32, device="cuda")
    eng.unit_vectors_dev(lat_d.data_ptr(), lon_d.data_ptr(), n, 1, xyz.data_ptr(), stream)
    ms = timed(lambda: eng.matrix_cells_dev(xyz.data_ptr(), n, 1, cells.data_ptr(), dist_scale=DIST_SCALE,
                                            time_scale=TS_CAR, stream=stream))
    out["matrix_10001_nodes"] = {"ms": ms, "cells_per_s": n * n / ms * 1e3, "GBs": n * n * 8 / ms / 1e6}
    del cells
    # host API (rt_matrix_build): 5,001 nodes, dense int32 planes back to pinned host memory
    n5 = 5001
    t_d = torch.empty((n5, n5), dtype=torch.int32).pin_memory()
    t_t = torch.empty((n5, n5), dtype=torch.int32).pin_memory()
    outb = {"dist": t_d.numpy(), "time": t_t.numpy()}
    t0 = time.perf_counter()
    for _ in range(3):
        eng.matrix_build(lat[:n5], lon[:n5], dist_scale=DIST_SCALE, time_scale=TS_CAR, out=outb)
    dt = (time.perf_counter() - t0) / 3
    out["matrix_host_api_5001_nodes"] = {"ms": dt * 1e3, "cells_per_s": n5 * n5 / dt, "d2h_GBs": n5 * n5 * 8 / dt / 1e9}
    # ETA: 1M pairs, synthetic ensemble of 200 depth-6 trees (complete trees, random splits)
    rng = np.random.default_rng(3456)
    trees, depth = 200, 6
    nodes = bytearray()
    offs = [0]
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
    blob = b"RTE1" + struct.pack("<IIf", trees, 12, 25.0) + struct.pack(f"<{len(offs)}I", *offs) + bytes(nodes)
    eng.eta_load_model(blob)
    m = 1_000_000
    cols = dict(d=torch.tensor(rng.uniform(0, 45000, m), device="cuda"),
                w=torch.tensor(rng.integers(0, 4, m).astype(np.uint8), device="cuda"),
                t=torch.tensor(rng.integers(0, 4, m).astype(np.uint8), device="cuda"),
                wd=torch.tensor(rng.integers(0, 7, m).astype(np.uint8), device="cuda"),
                h=torch.tensor(rng.integers(0, 24, m).astype(np.uint8), device="cuda"),
                a=torch.tensor(rng.uniform(20, 40, m).astype(np.float32), device="cuda"))
    res = torch.empty(m, dtype=torch.float32, device="cuda")
    ms = timed(lambda: eng.eta_score_dev(cols["d"].data_ptr(), cols["w"].data_ptr(), cols["t"].data_ptr(),
                                         cols["wd"].data_ptr(), cols["h"].data_ptr(), cols["a"].data_ptr(), m,
                                         res.data_ptr(), stream))
    out["eta_1M_pairs_200_trees_depth6"] = {"ms": ms, "pairs_per_s": m / ms * 1e3}
    # config 5: full solve of 5,000 stops / 50 vehicles, greedy start -> local optimum, batched iterations (SPEC §8.1)
    n = 5001
    lat, lon, demand, cap = synth_coords(5678, n - 1, 50)
    pr = eng.problem(n, 1)
    pr.set_coords(lat, lon)
    pr.build_matrix(DIST_SCALE, TS_CAR)
    _, t = pr.get_matrix(0)
    st = np.zeros((1, n, 4), np.int32)
    st[0, :, :3] = synth_windows(5678, t[0, 1:], t[1:, 0])
    st[0, :, 3] = demand
    pr.set_stops(st, [cap])
    t0 = time.perf_counter()
    trips, status = pr.greedy(demand[None, :].astype(float), [float(cap)], use_tw=True)
    t1 = time.perf_counter()
    pr.set_routes(trips)
    cost0 = int(pr.get_state(0)["cost"])
    t2 = time.perf_counter()
    moves, cost, _ = pr.local_search(batched=True)
    t3 = time.perf_counter()
    out["solve_5000_stops_50_vehicles"] = {"greedy_ms": (t1 - t0) * 1e3, "local_search_ms": (t3 - t2) * 1e3,
                                           "moves": int(moves[0]), "routes": len(trips[0]), "cost_greedy": cost0,
                                           "cost_local_optimum": int(cost[0]), "driver": "batched iterations (SPEC.md 8.1)"}
    pr.close()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="engine", choices=["engine", "reference"])
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--stops", type=int, default=1000)
    ap.add_argument("--vehicles", type=int, default=10)
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-extra", action="store_true", help="skip the secondary figures (other BASELINE configs)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank)
    else:
        run_engine(args, rank, world)


if __name__ == "__main__":
    main()
