"""Engine — Python face of the C ABI (include/routest.h).

One Engine per GPU and per process.  Methods without a `_dev` suffix take and return numpy arrays (host memory) and
do their own copies; `_dev` methods take raw device pointers (e.g. `torch.Tensor.data_ptr()`) and a CUDA stream
handle and never synchronise — torch supplies device memory, streams and NCCL, nothing else.
"""
from __future__ import annotations

import contextlib
import ctypes as C
import threading
import weakref

import numpy as np

from . import _native as N

# SPEC.md §3: speeds for the ORS profiles the reference names at utils.py:22-29 (it never asks ORS for durations).
PROFILE_SPEED_MPS = {
    "driving-car": 40.0 / 3.6,
    "driving-hgv": 35.0 / 3.6,
    "cycling-regular": 15.0 / 3.6,
    "cycling-road": 20.0 / 3.6,
    "foot-walking": 5.0 / 3.6,
}

_VEHICLE_PROFILE = {  # utils.py:22-29
    "car": "driving-car",
    "truck": "driving-hgv",
    "hgv": "driving-hgv",
    "motorcycle": "driving-car",
    "bike": "cycling-regular",
    "roadbike": "cycling-road",
    "foot": "foot-walking",
}

DEFAULT_DIST_SCALE = 10.0  # decimetres


def profile_for_vehicle(vehicle_type) -> str:
    """utils.py:19-29: lower-cased, stripped, default 'car' -> 'driving-car'."""
    v = (vehicle_type or "car").lower().strip()
    return _VEHICLE_PROFILE.get(v, "driving-car")


def time_scale_for_profile(profile: str, units_per_second: float = 1.0) -> float:
    """Time units per metre for a profile (SPEC.md §3)."""
    return units_per_second / PROFILE_SPEED_MPS.get(profile, PROFILE_SPEED_MPS["driving-car"])


class EngineError(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f"[rt error {code}] {message}")
        self.code = code
        self.message = message


def _f64(a):
    return a.ctypes.data_as(N.c_f64_p) if a is not None else None


def _i32(a):
    return a.ctypes.data_as(N.c_i32_p) if a is not None else None


class Engine:
    def __init__(self, device: int = 0):
        self._lib = N.lib()
        self._h = N.c_engine_p()
        rc = self._lib.rt_create(int(device), C.byref(self._h))
        if rc != N.RT_OK:
            msg = self._lib.rt_last_error(None)
            self._h = None
            raise EngineError(rc, (msg or b"rt_create failed").decode())
        self.device = int(device)
        self._problems = weakref.WeakSet()   # closed before the engine: a problem's handle points into it
        self._pool, self._pool_lock = {}, threading.Lock()   # idle single-instance problems by node count (lease_problem)

    # -- lifetime -----------------------------------------------------------------------------------------------
    def close(self):
        if getattr(self, "_h", None):
            with self._pool_lock:
                self._pool.clear()
            for pr in list(getattr(self, "_problems", ())):
                pr.close()
            self._lib.rt_destroy(self._h)
            self._h = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc: int):
        if rc != N.RT_OK:
            raise EngineError(rc, (self._lib.rt_last_error(self._h) or b"").decode())

    @property
    def handle(self):
        return self._h

    @property
    def stream(self) -> int:
        """The engine's own stream (cudaStream_t as int); `_dev` methods take a stream literally, 0 = default."""
        return int(self._lib.rt_engine_stream(self._h) or 0)

    # -- matrix (utils.py:94-109) -----------------------------------------------------------------------------------
    @staticmethod
    def matrix_pitch(n: int) -> int:
        return int(N.lib().rt_matrix_pitch(int(n)))

    def matrix_build(self, lat, lon, *, dist_scale: float = DEFAULT_DIST_SCALE, time_scale: float | None = None,
                     rows: tuple[int, int] | None = None, want=("dist", "time"), out=None):
        """Rows [rows[0], rows[1]) of the (n x n) matrix; node 0 is the depot.  Returns a dict of numpy arrays
        ('dist' int32, 'time' int32, 'metres' float64), each of shape (rows, n).  `out` may supply preallocated
        (e.g. pinned) arrays under the same keys."""
        lat = np.ascontiguousarray(lat, dtype=np.float64)
        lon = np.ascontiguousarray(lon, dtype=np.float64)
        if lat.ndim != 1 or lat.shape != lon.shape:
            raise ValueError("lat and lon must be 1-D arrays of equal length")
        n = lat.size
        r0, r1 = rows if rows is not None else (0, n)
        if time_scale is None:
            time_scale = time_scale_for_profile("driving-car")
        res = {}
        for key, dt in (("dist", np.int32), ("time", np.int32), ("metres", np.float64)):
            if key in want:
                a = out[key] if out is not None and key in out else np.empty((max(r1 - r0, 0), n), dtype=dt)
                if a.dtype != dt or a.shape != (r1 - r0, n) or not a.flags.c_contiguous:
                    raise ValueError(f"out[{key!r}] must be a C-contiguous {dt.__name__} array of shape {(r1 - r0, n)}")
                res[key] = a
        self._check(self._lib.rt_matrix_build(self._h, _f64(lat), _f64(lon), n, int(r0), int(r1), float(dist_scale),
                                              float(time_scale), _i32(res.get("dist")), _i32(res.get("time")),
                                              _f64(res.get("metres"))))
        return res

    def selftest_fast_sqrt(self, lo_bits: int, hi_bits: int) -> int:
        """Number of floats in the bit-pattern range whose straight-line sqrt differs from IEEE sqrtf (expect 0)."""
        bad = C.c_uint64(0)
        self._check(self._lib.rt_selftest_fast_sqrt(self._h, lo_bits, hi_bits, C.byref(bad)))
        return int(bad.value)

    def debug_set(self, name: str, value: int):
        """rt_debug_set: process-wide test/measurement knob ("strip_rows", "intra_kernel")."""
        self._check(self._lib.rt_debug_set(name.encode(), int(value)))

    def pipe_peak(self, mode: int = 0, iters: int = 20000):
        """Measured instruction-rate peak in warp instructions per second (rt_selftest_pipe_peak): mode 0 = ALU pipe,
        1 = FMA pipe (IMAD), 2 = ALU + IMAD interleaved, 3 = ALU + immediate-form FFMA interleaved (issue limit).  Returns (rate, elapsed_ms)."""
        rate, ms = C.c_double(0), C.c_double(0)
        self._check(self._lib.rt_selftest_pipe_peak(self._h, int(mode), int(iters), C.byref(rate), C.byref(ms)))
        return float(rate.value), float(ms.value)

    def unit_vectors_dev(self, lat_ptr: int, lon_ptr: int, n: int, batch: int, xyz_ptr: int, stream: int = 0):
        self._check(self._lib.rt_unit_vectors_dev(self._h, lat_ptr, lon_ptr, int(n), int(batch), xyz_ptr, stream or None))

    def matrix_cells_dev(self, xyz_ptr: int, n: int, batch: int, cells_ptr: int, *, rows=None,
                         dist_scale: float = DEFAULT_DIST_SCALE, time_scale: float | None = None, stream: int = 0):
        r0, r1 = rows if rows is not None else (0, n)
        if time_scale is None:
            time_scale = time_scale_for_profile("driving-car")
        self._check(self._lib.rt_matrix_cells_dev(self._h, xyz_ptr, int(n), int(batch), int(r0), int(r1),
                                                  float(dist_scale), float(time_scale), cells_ptr, stream or None))

    def matrix_cells_static_dev(self, lat_ptr: int, lon_ptr: int, n: int, cells_ptr: int, *,
                                dist_scale: float = DEFAULT_DIST_SCALE, time_scale: float | None = None, stream: int = 0):
        """rt_matrix_cells_static_dev: the device-resident matrix of one instance without any allocation — the unit
        vectors go through a static array inside the library instead of a caller buffer.  For processes whose GPU
        memory all belongs to a pool, or that keep cudaMalloc off the request path.  n <= static_max_nodes(); the
        array is shared per GPU, so keep calls on one stream; make the first call outside stream capture."""
        if time_scale is None:
            time_scale = time_scale_for_profile("driving-car")
        self._check(self._lib.rt_matrix_cells_static_dev(self._h, lat_ptr, lon_ptr, int(n), float(dist_scale),
                                                         float(time_scale), cells_ptr, stream or None))

    @staticmethod
    def static_max_nodes() -> int:
        return int(N.lib().rt_matrix_static_max_nodes())

    # -- ETA (ml.py:35-53) ------------------------------------------------------------------------------------------
    def eta_load_model(self, blob: bytes):
        buf = (C.c_char * len(blob)).from_buffer_copy(blob)
        self._check(self._lib.rt_eta_load_model(self._h, C.cast(buf, C.c_void_p), len(blob)))

    def eta_model_info(self):
        nt, nn = C.c_int32(0), C.c_int32(0)
        self._check(self._lib.rt_eta_model_info(self._h, C.byref(nt), C.byref(nn)))
        return int(nt.value), int(nn.value)

    def _eta_cols(self, distance_m, weather, traffic, weekday, hour, age):
        d = np.ascontiguousarray(distance_m, dtype=np.float64)
        n = d.size
        cols = [np.ascontiguousarray(a, dtype=np.uint8) for a in (weather, traffic, weekday, hour)]
        a = np.ascontiguousarray(age, dtype=np.float32)
        if any(c.size != n for c in cols) or a.size != n:
            raise ValueError("all ETA input columns must have the same length")
        return d, cols, a, n

    def eta_score(self, distance_m, weather, traffic, weekday, hour, age):
        """Minutes (float32[n]) for n stop pairs; columns as in include/routest.h (rt_eta_score)."""
        d, cols, a, n = self._eta_cols(distance_m, weather, traffic, weekday, hour, age)
        out = np.zeros(n, np.float32)
        fp = C.POINTER(C.c_float)
        self._check(self._lib.rt_eta_score(self._h, _f64(d), *[c.ctypes.data_as(N.c_u8_p) for c in cols],
                                           a.ctypes.data_as(fp), n, out.ctypes.data_as(fp)))
        return out

    def eta_features(self, distance_m, weather, traffic, weekday, hour, age):
        """The float32 [n,12] rows model.predict is handed (ml.py:35-48 column order)."""
        d, cols, a, n = self._eta_cols(distance_m, weather, traffic, weekday, hour, age)
        out = np.zeros((n, 12), np.float32)
        fp = C.POINTER(C.c_float)
        self._check(self._lib.rt_eta_features(self._h, _f64(d), *[c.ctypes.data_as(N.c_u8_p) for c in cols],
                                              a.ctypes.data_as(fp), n, out.ctypes.data_as(fp)))
        return out

    def eta_score_dev(self, dist_ptr, weather_ptr, traffic_ptr, weekday_ptr, hour_ptr, age_ptr, n, out_ptr, stream=0):
        self._check(self._lib.rt_eta_score_dev(self._h, dist_ptr, weather_ptr, traffic_ptr, weekday_ptr, hour_ptr,
                                               age_ptr, int(n), out_ptr, stream or None))

    # -- greedy on a caller-supplied matrix (utils.py:111-139) --------------------------------------------------------
    def greedy_trips(self, D, demand, cap: float, max_dist: float):
        """Trips over an (n x n) matrix D held by the caller (floats allowed, asymmetric allowed), node 0 = depot.
        Returns (trips, infeasible): trips is a list of lists of matrix indices without the depot; infeasible is
        True where the reference would loop forever (SPEC.md §4)."""
        D = np.ascontiguousarray(D, dtype=np.float64)
        n = D.shape[0]
        if D.shape != (n, n):
            raise ValueError("D must be square")
        demand = np.ascontiguousarray(demand, dtype=np.float64)
        if demand.shape != (n,):
            raise ValueError("demand must have one entry per node (entry 0 is ignored)")
        order = np.zeros(max(n - 1, 1), np.int32)
        off = np.zeros(n, np.int32)
        nt = C.c_int32(0)
        rc = self._lib.rt_greedy_trips(self._h, _f64(D), n, _f64(demand), float(cap), float(max_dist), _i32(order),
                                       _i32(off), C.byref(nt))
        if rc not in (N.RT_OK, N.RT_ERR_INFEASIBLE_STOP):
            self._check(rc)
        trips = [order[off[k]:off[k + 1]].tolist() for k in range(nt.value)]
        return trips, rc == N.RT_ERR_INFEASIBLE_STOP

    def problem(self, n: int, batch: int = 1) -> "Problem":
        return Problem(self, n, batch)

    POOL_MAX_NODES = 2048      # larger problems are created and destroyed per use (their matrix alone is > 32 MB)
    POOL_MAX_IDLE_PER_SIZE = 4

    @contextlib.contextmanager
    def lease_problem(self, n: int):
        """A single-instance problem of n nodes for the duration of one request.  rt_problem_create / _destroy are some
        forty cudaMalloc / cudaFree calls — 0.3-0.4 ms, most of a small request — so idle problems are kept per node
        count and handed out again (the reference's handlers run on arbitrary Flask worker threads, app.py:8: concurrent
        requests each get their own problem; a size seen for the first time is created on the spot).  The caller
        re-installs coordinates / matrix / stops / routes as it goes; nothing of the previous request is read."""
        pr = None
        if n <= self.POOL_MAX_NODES:
            with self._pool_lock:
                idle = self._pool.get(n)
                if idle:
                    pr = idle.pop()
        if pr is None:
            pr = Problem(self, n, 1)
        try:
            yield pr
        except BaseException:
            pr.close()   # whatever went wrong, do not hand this one out again
            raise
        else:
            keep = False
            if n <= self.POOL_MAX_NODES and getattr(pr, "_p", None):
                with self._pool_lock:
                    idle = self._pool.setdefault(n, [])
                    if len(idle) < self.POOL_MAX_IDLE_PER_SIZE:
                        idle.append(pr)
                        keep = True
            if not keep:
                pr.close()


NO_MOVE = 0xFFFFFFFFFFFFFFFF
FAM_2OPT, FAM_RELOCATE, FAM_SWAP = 0, 1, 2
ALL_FAMILIES = 7


def key_unpack(k: int):
    """(delta, family, p, q) of a packed move key (SPEC.md §7)."""
    k = int(k)
    return ((k >> 32) - 0x80000000, (k >> 30) & 3, (k >> 15) & 0x7FFF, k & 0x7FFF)


def _u64(a):
    return a.ctypes.data_as(N.c_u64_p)


def _i64(a):
    return a.ctypes.data_as(N.c_i64_p) if a is not None else None


class Problem:
    """A batch of device-resident instances (include/routest.h: rt_problem).  Arrays are instance-major."""

    def __init__(self, engine: Engine, n: int, batch: int = 1):
        self.engine = engine
        self._lib = engine._lib
        self._p = N.c_problem_p()
        engine._check(self._lib.rt_problem_create(engine.handle, int(n), int(batch), C.byref(self._p)))
        self.n, self.batch = int(n), int(batch)
        self.pstride = int(self._lib.rt_problem_position_stride(self._p))
        engine._problems.add(self)

    def close(self):
        if getattr(self, "_p", None):
            self._lib.rt_problem_destroy(self._p)
            self._p = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def handle(self):
        return self._p

    def _check(self, rc):
        self.engine._check(rc)

    def _arr(self, a, dtype, shape):
        a = np.ascontiguousarray(a, dtype=dtype)
        if a.shape != shape:
            if a.size == int(np.prod(shape)):
                a = a.reshape(shape)
            else:
                raise ValueError(f"expected shape {shape}, got {a.shape}")
        return a

    # -- inputs
    def set_coords(self, lat, lon):
        lat = self._arr(lat, np.float64, (self.batch, self.n))
        lon = self._arr(lon, np.float64, (self.batch, self.n))
        self._check(self._lib.rt_problem_set_coords(self._p, _f64(lat), _f64(lon)))

    def set_stops(self, stops, cap, max_dist=None):
        stops = self._arr(stops, np.int32, (self.batch, self.n, 4))
        cap = self._arr(np.broadcast_to(np.asarray(cap), (self.batch,)), np.int32, (self.batch,))
        md = np.full(self.batch, 2 ** 62, np.int64) if max_dist is None else \
            self._arr(np.broadcast_to(np.asarray(max_dist), (self.batch,)), np.int64, (self.batch,))
        self._check(self._lib.rt_problem_set_stops(self._p, _i32(stops), _i32(cap), _i64(md)))

    def build_matrix(self, dist_scale: float = DEFAULT_DIST_SCALE, time_scale: float | None = None):
        if time_scale is None:
            time_scale = time_scale_for_profile("driving-car")
        self._check(self._lib.rt_problem_build_matrix(self._p, float(dist_scale), float(time_scale)))

    def set_matrix(self, instance: int, dist, time):
        dist = self._arr(dist, np.int32, (self.n, self.n))
        time = self._arr(time, np.int32, (self.n, self.n))
        self._check(self._lib.rt_problem_set_matrix(self._p, int(instance), _i32(dist), _i32(time)))

    def get_matrix(self, instance: int = 0):
        d = np.empty((self.n, self.n), np.int32)
        t = np.empty((self.n, self.n), np.int32)
        self._check(self._lib.rt_problem_get_matrix(self._p, int(instance), _i32(d), _i32(t)))
        return d, t

    # -- construction (utils.py:111-139)
    def greedy(self, demand, cap, max_dist=9e12, use_tw: bool = False):
        """Returns (trips per instance, status per instance)."""
        demand = self._arr(demand, np.float64, (self.batch, self.n))
        cap = self._arr(np.broadcast_to(np.asarray(cap, dtype=np.float64), (self.batch,)), np.float64, (self.batch,))
        md = self._arr(np.broadcast_to(np.asarray(max_dist, dtype=np.float64), (self.batch,)), np.float64, (self.batch,))
        order = np.zeros((self.batch, self.n - 1), np.int32)
        off = np.zeros((self.batch, self.n), np.int32)
        nt = np.zeros(self.batch, np.int32)
        status = np.zeros(self.batch, np.int32)
        self._check(self._lib.rt_problem_greedy(self._p, _f64(demand), _f64(cap), _f64(md), int(bool(use_tw)),
                                                _i32(order), _i32(off), _i32(nt), _i32(status)))
        trips = [[order[b, off[b, k]:off[b, k + 1]].tolist() for k in range(nt[b])] for b in range(self.batch)]
        return trips, status

    # -- routes and state (SPEC.md §6)
    def set_routes(self, trips):
        """trips: per instance, a list of lists of customer ids."""
        order = np.zeros((self.batch, self.n - 1), np.int32)
        off = np.zeros((self.batch, self.n), np.int32)
        nt = np.zeros(self.batch, np.int32)
        for b, tr in enumerate(trips):
            flat = [c for t in tr for c in t]
            order[b, :len(flat)] = flat
            nt[b] = len(tr)
            off[b, 1:len(tr) + 1] = np.cumsum([len(t) for t in tr])
        self._check(self._lib.rt_problem_set_routes(self._p, _i32(order), _i32(off), _i32(nt)))

    def get_routes(self):
        order = np.zeros((self.batch, self.n - 1), np.int32)
        off = np.zeros((self.batch, self.n), np.int32)
        nt = np.zeros(self.batch, np.int32)
        self._check(self._lib.rt_problem_get_routes(self._p, _i32(order), _i32(off), _i32(nt)))
        return [[order[b, off[b, k]:off[b, k + 1]].tolist() for k in range(nt[b])] for b in range(self.batch)]

    def get_state(self, instance: int = 0):
        ps, n = self.pstride, self.n
        npos = C.c_int32(0)
        node, route, B, Z = (np.zeros(ps, np.int32) for _ in range(4))
        rload, rdist = np.zeros(n, np.int64), np.zeros(n, np.int64)
        cost, feas = C.c_int64(0), C.c_int32(0)
        self._check(self._lib.rt_problem_get_state(self._p, int(instance), C.byref(npos), _i32(node), _i32(route),
                                                   _i32(B), _i32(Z), _i64(rload), _i64(rdist), C.byref(cost),
                                                   C.byref(feas)))
        P = npos.value
        R = P - n
        return dict(P=P, node=node[:P], route=route[:P], B=B[:P], Z=Z[:P], rload=rload[:R], rdist=rdist[:R],
                    cost=int(cost.value), feasible=bool(feas.value))

    def get_position_matrix(self, instance: int, P: int):
        """The route-order matrix the move kernels read: (dist, time), each int32 [P, P]."""
        d = np.empty((P, P), np.int32)
        t = np.empty((P, P), np.int32)
        self._check(self._lib.rt_problem_get_position_matrix(self._p, int(instance), _i32(d), _i32(t)))
        return d, t

    # -- moves (SPEC.md §7-§8)
    def eval_moves(self, families: int = ALL_FAMILIES, shard=(0, 1)):
        keys = np.zeros(self.batch, np.uint64)
        self._check(self._lib.rt_problem_eval_moves(self._p, int(families), int(shard[0]), int(shard[1]), _u64(keys)))
        return keys

    def eval_dump(self, instance: int, family: int, P: int):
        delta = np.zeros((P, P), np.int32)
        feas = np.zeros((P, P), np.uint8)
        self._check(self._lib.rt_problem_eval_dump(self._p, int(instance), int(family), _i32(delta),
                                                   feas.ctypes.data_as(N.c_u8_p)))
        return delta, feas

    def apply_moves(self, keys):
        keys = self._arr(keys, np.uint64, (self.batch,))
        self._check(self._lib.rt_problem_apply_moves(self._p, _u64(keys)))

    def local_search(self, families: int = ALL_FAMILIES, max_moves: int = 1 << 30, keys_capacity: int = 0,
                     full: bool = False, batched: bool = False):
        """Best-improvement search to a local optimum.  full=True re-scores the whole neighbourhood every iteration
        (rt_problem_local_search_full) instead of only what the last move touched; the result is the same.
        batched=True applies a set of independent improving moves per iteration (SPEC.md §8.1,
        rt_problem_local_search_batched): max_moves then bounds the iterations, and the local optimum reached is in
        general a different one."""
        moves = np.zeros(self.batch, np.int32)
        cost = np.zeros(self.batch, np.int64)
        keys = np.zeros((self.batch, max(keys_capacity, 1)), np.uint64)
        fn = (self._lib.rt_problem_local_search_batched if batched else
              self._lib.rt_problem_local_search_full if full else self._lib.rt_problem_local_search)
        self._check(fn(self._p, int(families), int(min(max_moves, 2 ** 31 - 1)), _i32(moves), _i64(cost),
                       _u64(keys) if keys_capacity else None, int(keys_capacity)))
        return moves, cost, [keys[b, :min(moves[b], keys_capacity)].copy() for b in range(self.batch)]

    def plan(self, lat, lon, stops, cap, max_dist, node, npos, *, dist_scale=DEFAULT_DIST_SCALE, time_scale=None,
             families: int = ALL_FAMILIES, out=None):
        """One pass of the hot path from host arrays to the best move of every instance (rt_problem_plan).  The
        arrays are used as given when they already have the right dtype/shape (so pinned buffers stay pinned)."""
        if time_scale is None:
            time_scale = time_scale_for_profile("driving-car")
        lat = self._arr(lat, np.float64, (self.batch, self.n))
        lon = self._arr(lon, np.float64, (self.batch, self.n))
        stops = self._arr(stops, np.int32, (self.batch, self.n, 4))
        cap = self._arr(cap, np.int32, (self.batch,))
        max_dist = self._arr(max_dist, np.int64, (self.batch,))
        node = self._arr(node, np.int32, (self.batch, self.pstride))
        npos = self._arr(npos, np.int32, (self.batch,))
        keys = out if out is not None else np.zeros(self.batch, np.uint64)
        self._check(self._lib.rt_problem_plan(self._p, _f64(lat), _f64(lon), _i32(stops), _i32(cap), _i64(max_dist),
                                              _i32(node), _i32(npos), float(dist_scale), float(time_scale),
                                              int(families), _u64(keys)))
        return keys

    def plan_call(self, lat, lon, stops, cap, max_dist, node, npos, out, *, dist_scale=DEFAULT_DIST_SCALE,
                  time_scale=None, families: int = ALL_FAMILIES):
        """`plan` for callers that reuse the same host buffers request after request: the arrays are checked and
        their addresses taken once; the returned function makes exactly one C call per invocation and returns
        `out`.  The buffers must keep their dtype, shape and address (fill them in place, e.g. pinned arrays)."""
        if time_scale is None:
            time_scale = time_scale_for_profile("driving-car")
        spec = ((lat, np.float64, (self.batch, self.n)), (lon, np.float64, (self.batch, self.n)),
                (stops, np.int32, (self.batch, self.n, 4)), (cap, np.int32, (self.batch,)),
                (max_dist, np.int64, (self.batch,)), (node, np.int32, (self.batch, self.pstride)),
                (npos, np.int32, (self.batch,)), (out, np.uint64, (self.batch,)))
        for a, dt, shape in spec:
            if not isinstance(a, np.ndarray) or a.dtype != dt or a.shape != shape or not a.flags.c_contiguous:
                raise ValueError(f"plan_call needs C-contiguous {np.dtype(dt).name} arrays of shape {shape}")
        fn, check, handle = self._lib.rt_problem_plan, self._check, self._p
        args = (handle, _f64(lat), _f64(lon), _i32(stops), _i32(cap), _i64(max_dist), _i32(node), _i32(npos),
                C.c_double(dist_scale), C.c_double(time_scale), C.c_int(families), _u64(out))
        keep = (lat, lon, stops, cap, max_dist, node, npos, out)

        def call(_keep=keep):
            check(fn(*args))
            return out
        return call

    def build_matrix_dev(self, dist_scale=DEFAULT_DIST_SCALE, time_scale=None, stream: int = 0):
        if time_scale is None:
            time_scale = time_scale_for_profile("driving-car")
        self._check(self._lib.rt_problem_build_matrix_dev(self._p, float(dist_scale), float(time_scale), stream or None))

    def refresh_dev(self, what: int = 3, stream: int = 0):
        """what: bit 0 = route-order matrix, bit 1 = route state tables."""
        self._check(self._lib.rt_problem_refresh_dev(self._p, int(what), stream or None))

    def plan_pipeline_calls(self, lat, lon, stops, cap, max_dist, node, npos, out, *, dist_scale=DEFAULT_DIST_SCALE,
                            time_scale=None, families: int = ALL_FAMILIES):
        """(begin, end) for callers that keep requests in flight (rt_problem_plan_begin / rt_problem_plan_end): `begin()`
        enqueues this problem's uploads and kernels and returns at once, `end()` waits for them and returns `out`.
        With two problems of one engine used in turn the upload of one request overlaps the kernels of the other.
        Same buffer rules as plan_call; the inputs must stay unchanged between begin() and end()."""
        if time_scale is None:
            time_scale = time_scale_for_profile("driving-car")
        spec = ((lat, np.float64, (self.batch, self.n)), (lon, np.float64, (self.batch, self.n)),
                (stops, np.int32, (self.batch, self.n, 4)), (cap, np.int32, (self.batch,)),
                (max_dist, np.int64, (self.batch,)), (node, np.int32, (self.batch, self.pstride)),
                (npos, np.int32, (self.batch,)), (out, np.uint64, (self.batch,)))
        for a, dt, shape in spec:
            if not isinstance(a, np.ndarray) or a.dtype != dt or a.shape != shape or not a.flags.c_contiguous:
                raise ValueError(f"plan_pipeline_calls needs C-contiguous {np.dtype(dt).name} arrays of shape {shape}")
        fb, fe, check, handle = self._lib.rt_problem_plan_begin, self._lib.rt_problem_plan_end, self._check, self._p
        args = (handle, _f64(lat), _f64(lon), _i32(stops), _i32(cap), _i64(max_dist), _i32(node), _i32(npos),
                C.c_double(dist_scale), C.c_double(time_scale), C.c_int(families))
        out_p = _u64(out)
        keep = (lat, lon, stops, cap, max_dist, node, npos, out)

        def begin(_keep=keep):
            check(fb(*args))

        def end():
            check(fe(handle, out_p))
            return out
        return begin, end

    def plan_upload_call(self, lat, lon, stops, cap, max_dist, node, npos, keys_ptr: int, *,
                         dist_scale=DEFAULT_DIST_SCALE, time_scale=None, families: int = ALL_FAMILIES, stream: int = 0):
        """rt_problem_plan_upload_dev as a reusable call (same buffer rules as plan_call): host inputs go up on
        `stream`, this replica's shard is scored, the signed-order keys stay on the device at keys_ptr (int64[batch]);
        nothing is synchronised.  What a rank of a multi-GPU search does per request before the MIN all-reduce."""
        if time_scale is None:
            time_scale = time_scale_for_profile("driving-car")
        spec = ((lat, np.float64, (self.batch, self.n)), (lon, np.float64, (self.batch, self.n)),
                (stops, np.int32, (self.batch, self.n, 4)), (cap, np.int32, (self.batch,)),
                (max_dist, np.int64, (self.batch,)), (node, np.int32, (self.batch, self.pstride)),
                (npos, np.int32, (self.batch,)))
        for a, dt, shape in spec:
            if not isinstance(a, np.ndarray) or a.dtype != dt or a.shape != shape or not a.flags.c_contiguous:
                raise ValueError(f"plan_upload_call needs C-contiguous {np.dtype(dt).name} arrays of shape {shape}")
        fn, check = self._lib.rt_problem_plan_upload_dev, self._check
        args = (self._p, _f64(lat), _f64(lon), _i32(stops), _i32(cap), _i64(max_dist), _i32(node), _i32(npos),
                C.c_double(dist_scale), C.c_double(time_scale), C.c_int(families), C.c_void_p(keys_ptr),
                C.c_void_p(stream or None))
        keep = (lat, lon, stops, cap, max_dist, node, npos)

        def call(_keep=keep):
            check(fn(*args))
        return call

    def plan_dev(self, keys_ptr: int, *, dist_scale=DEFAULT_DIST_SCALE, time_scale=None, families: int = ALL_FAMILIES,
                 stream: int = 0):
        if time_scale is None:
            time_scale = time_scale_for_profile("driving-car")
        self._check(self._lib.rt_problem_plan_dev(self._p, float(dist_scale), float(time_scale), int(families),
                                                  keys_ptr, stream or None))

    def set_shard(self, index: int, count: int):
        """This replica evaluates part `index` of `count` of the candidate space (multi-GPU search)."""
        self._check(self._lib.rt_problem_set_shard(self._p, int(index), int(count)))

    def eval_moves_dev(self, keys_ptr: int, families: int = ALL_FAMILIES, shard=(0, 1), stream: int = 0):
        self._check(self._lib.rt_problem_eval_moves_dev(self._p, int(families), int(shard[0]), int(shard[1]), keys_ptr,
                                                        stream or None))

    def apply_moves_dev(self, keys_ptr: int, improving_only: bool = True, stream: int = 0):
        self._check(self._lib.rt_problem_apply_moves_dev(self._p, keys_ptr, int(bool(improving_only)), stream or None))

    # -- batched iterations over replicas (SPEC.md §8.1): table per rank -> MIN all-reduce -> same batch everywhere
    def table_entries(self) -> int:
        n = int(self._lib.rt_problem_table_entries(self._p))
        if n <= 0:
            raise EngineError(N.RT_ERR_STATE, "no routes installed, or more than 1024 routes in an instance")
        return n

    def eval_table_dev(self, table_ptr: int, families: int = ALL_FAMILIES, shard=(0, 1), stream: int = 0):
        self._check(self._lib.rt_problem_eval_table_dev(self._p, int(families), int(shard[0]), int(shard[1]), table_ptr,
                                                        stream or None))

    def apply_batch_dev(self, table_ptr: int, stream: int = 0):
        self._check(self._lib.rt_problem_apply_batch_dev(self._p, table_ptr, stream or None))

    # -- peer-memory exchange (multi-GPU search without a collective call per iteration)
    def exchange_export(self) -> bytes:
        buf = C.create_string_buffer(int(self._lib.rt_problem_exchange_handle_bytes()))
        self._check(self._lib.rt_problem_exchange_export(self._p, C.cast(buf, C.c_void_p)))
        return buf.raw

    def exchange_connect(self, handles, rank: int, world: int):
        blob = b"".join(handles)
        buf = C.create_string_buffer(blob, len(blob))
        self._check(self._lib.rt_problem_exchange_connect(self._p, C.cast(buf, C.c_void_p), int(rank), int(world)))

    def eval_moves_xchg_dev(self, families: int = ALL_FAMILIES, stream: int = 0):
        self._check(self._lib.rt_problem_eval_moves_xchg_dev(self._p, int(families), stream or None))

    def apply_moves_xchg_dev(self, improving_only: bool = True, stream: int = 0):
        self._check(self._lib.rt_problem_apply_moves_xchg_dev(self._p, int(bool(improving_only)), stream or None))

    def collect_keys_xchg_dev(self, keys_ptr: int, stream: int = 0):
        self._check(self._lib.rt_problem_collect_keys_xchg_dev(self._p, keys_ptr, stream or None))

    def table_exchange_export(self, world: int) -> bytes:
        buf = C.create_string_buffer(int(self._lib.rt_problem_exchange_handle_bytes()))
        self._check(self._lib.rt_problem_table_exchange_export(self._p, int(world), C.cast(buf, C.c_void_p)))
        return buf.raw

    def table_exchange_connect(self, handles, rank: int, world: int):
        blob = b"".join(handles)
        buf = C.create_string_buffer(blob, len(blob))
        self._check(self._lib.rt_problem_table_exchange_connect(self._p, C.cast(buf, C.c_void_p), int(rank), int(world)))

    def eval_table_xchg_dev(self, table_ptr: int, families: int = ALL_FAMILIES, stream: int = 0):
        self._check(self._lib.rt_problem_eval_table_xchg_dev(self._p, int(families), table_ptr, stream or None))

    def apply_batch_xchg_dev(self, table_ptr: int, stream: int = 0):
        self._check(self._lib.rt_problem_apply_batch_xchg_dev(self._p, table_ptr, stream or None))

    def search_status(self):
        done = np.zeros(self.batch, np.int32)
        moves = np.zeros(self.batch, np.int32)
        cost = np.zeros(self.batch, np.int64)
        self._check(self._lib.rt_problem_search_status(self._p, _i32(done), _i32(moves), _i64(cost)))
        return done, moves, cost
