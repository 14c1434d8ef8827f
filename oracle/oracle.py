"""ctypes front end of oracle/liboracle.so (see routest_oracle.c for what each function restates).

TEST INFRASTRUCTURE ONLY: the product (routest_b200/) never imports this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

OK = 0
ERR_INFEASIBLE_STOP = 5


def build(force: bool = False) -> str:
    so = os.path.join(_HERE, "liboracle.so")
    srcs = [os.path.join(_HERE, f) for f in ("routest_oracle.c", "routest_moves.c", "Makefile")]
    if force or not os.path.exists(so) or os.path.getmtime(so) < max(os.path.getmtime(f) for f in srcs):
        subprocess.check_call(["make", "-C", _HERE, "-B", "liboracle.so"], stdout=subprocess.DEVNULL)
    return so


def lib() -> C.CDLL:
    global _LIB
    if _LIB is None:
        _LIB = C.CDLL(build())
        _LIB.or_num_threads.restype = C.c_int
        _LIB.or_greedy_trips.restype = C.c_int
        _LIB.or_p2p_check.restype = C.c_int
        _LIB.or_p2p_check.argtypes = [C.c_double] * 4
        _LIB.or_build_positions.restype = C.c_int
        _LIB.or_route_state.restype = C.c_int
        _LIB.or_best_move.restype = C.c_uint64
        _LIB.or_local_search.restype = C.c_int
        _LIB.or_select_batch.restype = C.c_int
        _LIB.or_local_search_batched.restype = C.c_int
    return _LIB


def _p(a, t):
    return a.ctypes.data_as(C.POINTER(t)) if a is not None else None


def num_threads() -> int:
    return int(lib().or_num_threads())


def sincos(t):
    t = np.ascontiguousarray(t, dtype=np.float64)
    s = np.empty_like(t)
    c = np.empty_like(t)
    lib().or_sincos_vec(_p(t, C.c_double), C.c_int(t.size), _p(s, C.c_double), _p(c, C.c_double))
    return s, c


def unit_vectors(lat, lon):
    lat = np.ascontiguousarray(lat, dtype=np.float64)
    lon = np.ascontiguousarray(lon, dtype=np.float64)
    n = lat.size
    xyz = np.empty((3, n), dtype=np.float64)
    lib().or_unit_vectors(_p(lat, C.c_double), _p(lon, C.c_double), C.c_int(n), _p(xyz, C.c_double))
    return xyz


def matrix_rows(xyz, r0, r1, dist_scale, time_scale, want=("dist", "time"), threads=1):
    """Rows [r0,r1) of the matrix of SPEC §2.  Returns a dict with the requested planes."""
    xyz = np.ascontiguousarray(xyz, dtype=np.float64)
    n = xyz.shape[1]
    rows = r1 - r0
    out = {}
    d = np.empty((rows, n), np.int32) if "dist" in want else None
    t = np.empty((rows, n), np.int32) if "time" in want else None
    m = np.empty((rows, n), np.float64) if "metres" in want else None
    lib().or_matrix_rows(_p(xyz, C.c_double), C.c_int(n), C.c_int(r0), C.c_int(r1), C.c_double(dist_scale),
                         C.c_double(time_scale), _p(d, C.c_int32), _p(t, C.c_int32), _p(m, C.c_double),
                         C.c_int(threads))
    if d is not None:
        out["dist"] = d
    if t is not None:
        out["time"] = t
    if m is not None:
        out["metres"] = m
    return out


def haversine_libm(lat, lon):
    lat = np.ascontiguousarray(lat, dtype=np.float64)
    lon = np.ascontiguousarray(lon, dtype=np.float64)
    n = lat.size
    m = np.empty((n, n), np.float64)
    lib().or_haversine_libm(_p(lat, C.c_double), _p(lon, C.c_double), C.c_int(n), _p(m, C.c_double))
    return m


def greedy_trips(D, demand, cap, max_dist, T=None, tw=None):
    """SPEC §4.  D: n x n (any numeric dtype, converted to FP64 exactly as Python would).  demand: length n
    (index 0 ignored).  Returns (rc, trips) with trips a list of lists of matrix indices (no depot entries)."""
    D = np.ascontiguousarray(D, dtype=np.float64)
    n = D.shape[0]
    demand = np.ascontiguousarray(demand, dtype=np.float64)
    assert demand.size == n
    if tw is not None:
        tw = np.ascontiguousarray(tw, dtype=np.int32)
        T = np.ascontiguousarray(T, dtype=np.int32)
        assert tw.shape == (n, 3) and T.shape == (n, n)
    order = np.zeros(max(n - 1, 1), np.int32)
    off = np.zeros(n + 1, np.int32)
    nt = C.c_int32(0)
    rc = lib().or_greedy_trips(_p(D, C.c_double), _p(T, C.c_int32), _p(tw, C.c_int32), C.c_int(n),
                               _p(demand, C.c_double), C.c_double(cap), C.c_double(max_dist),
                               _p(order, C.c_int32), _p(off, C.c_int32), C.byref(nt))
    trips = [order[off[k]:off[k + 1]].tolist() for k in range(nt.value)]
    return int(rc), trips


def p2p_check(payload, cap, dist_m, max_dist) -> int:
    return int(lib().or_p2p_check(float(payload), float(cap), float(dist_m), float(max_dist)))


# ----------------------------------------------------------------------------------------------- SPEC §6-§8

NO_MOVE = 0xFFFFFFFFFFFFFFFF
FAM_2OPT, FAM_RELOCATE, FAM_SWAP = 0, 1, 2
ALL_FAMILIES = 7


def key_unpack(k: int):
    """(delta, family, p, q) of a packed key (SPEC §7)."""
    k = int(k)
    return ((k >> 32) - 0x80000000, (k >> 30) & 3, (k >> 15) & 0x7FFF, k & 0x7FFF)


def key_pack(delta: int, family: int, p: int, q: int) -> int:
    return ((delta + 0x80000000) << 32) | (family << 30) | (p << 15) | q


def build_positions(trips, n):
    """trips: list of lists of customer ids -> node[] position array (SPEC §6)."""
    order = np.array([i for t in trips for i in t], dtype=np.int32)
    off = np.zeros(len(trips) + 1, np.int32)
    off[1:] = np.cumsum([len(t) for t in trips])
    node = np.zeros(n + len(trips), np.int32)
    P = lib().or_build_positions(_p(order, C.c_int32), _p(off, C.c_int32), C.c_int(len(trips)), _p(node, C.c_int32))
    assert P == node.size
    return node


def positions_to_trips(node):
    trips, cur = [], []
    for v in list(node)[1:]:
        if v == 0:
            trips.append(cur)
            cur = []
        else:
            cur.append(int(v))
    return trips


class Instance:
    """Dense int32 D/T planes + stops {ready,due,service,demand} + limits, as the oracle functions take them."""

    def __init__(self, D, T, stops, cap, max_dist=2 ** 62):
        self.D = np.ascontiguousarray(D, dtype=np.int32)
        self.T = np.ascontiguousarray(T, dtype=np.int32)
        self.stops = np.ascontiguousarray(stops, dtype=np.int32)
        self.n = self.D.shape[0]
        assert self.D.shape == (self.n, self.n) and self.T.shape == (self.n, self.n) and self.stops.shape == (self.n, 4)
        self.cap = int(cap)
        self.max_dist = int(max_dist)

    def _args(self):
        return (C.c_int(self.n), _p(self.D, C.c_int32), _p(self.T, C.c_int32), _p(self.stops, C.c_int32),
                C.c_int32(self.cap), C.c_int64(self.max_dist))


def route_state(inst: Instance, node):
    node = np.ascontiguousarray(node, dtype=np.int32)
    P = node.size
    R = P - inst.n
    route = np.zeros(P, np.int32)
    B = np.zeros(P, np.int32)
    Z = np.zeros(P, np.int32)
    rload = np.zeros(P, np.int64)
    rdist = np.zeros(P, np.int64)
    cost = C.c_int64(0)
    feas = lib().or_route_state(*inst._args(), _p(node, C.c_int32), C.c_int(P), _p(route, C.c_int32), _p(B, C.c_int32),
                                _p(Z, C.c_int32), _p(rload, C.c_int64), _p(rdist, C.c_int64), C.byref(cost))
    return dict(feasible=bool(feas), route=route, B=B, Z=Z, rload=rload[:R], rdist=rdist[:R], cost=int(cost.value))


def _eval(fn, inst, node, family):
    node = np.ascontiguousarray(node, dtype=np.int32)
    P = node.size
    delta = np.zeros((P, P), np.int32)
    feas = np.zeros((P, P), np.uint8)
    fn(*inst._args(), _p(node, C.c_int32), C.c_int(P), C.c_int(family), _p(delta, C.c_int32), _p(feas, C.c_uint8))
    return delta, feas


def eval_bruteforce(inst, node, family):
    """(delta[P,P], feas[P,P]) by apply-and-recompute; feas: 0 infeasible, 1 feasible, 2 not a candidate."""
    return _eval(lib().or_eval_bruteforce, inst, node, family)


def eval_fast(inst, node, family):
    """Same through the O(1)/sweep algebra of SPEC §7 (what the engine implements)."""
    return _eval(lib().or_eval_fast, inst, node, family)


def best_move(inst, node, families=ALL_FAMILIES, shard=None, threads=1) -> int:
    node = np.ascontiguousarray(node, dtype=np.int32)
    P = node.size
    p0, p1 = shard if shard is not None else (0, P)
    return int(lib().or_best_move(*inst._args(), _p(node, C.c_int32), C.c_int(P), C.c_int(families), C.c_int(p0),
                                  C.c_int(p1), C.c_int(threads), None))


def best_from_dump(delta, feas, family):
    """Minimum key over a dumped family (for cross-checking best_move against the dumps)."""
    ok = feas == 1
    if not ok.any():
        return NO_MOVE
    P = delta.shape[0]
    p, q = np.nonzero(ok)
    keys = ((delta[p, q].astype(np.int64) + 0x80000000).astype(np.uint64) << np.uint64(32)) | \
        (np.uint64(family) << np.uint64(30)) | (p.astype(np.uint64) << np.uint64(15)) | q.astype(np.uint64)
    return int(keys.min())


def count_candidates(node):
    node = np.ascontiguousarray(node, dtype=np.int32)
    counts = np.zeros(3, np.int64)
    lib().or_count_candidates(_p(node, C.c_int32), C.c_int(node.size), _p(counts, C.c_int64))
    return counts


def apply_move(node, family, p, q):
    node = np.array(node, dtype=np.int32)
    lib().or_apply_move(_p(node, C.c_int32), C.c_int(family), C.c_int(p), C.c_int(q))
    return node


def local_search(inst, node, families=ALL_FAMILIES, max_moves=1 << 30, threads=1):
    """Returns (final node[], applied keys, final cost)."""
    node = np.array(node, dtype=np.int32)
    cap_keys = min(max_moves, 1 << 22)
    keys = np.zeros(cap_keys, np.uint64)
    cost = C.c_int64(0)
    m = lib().or_local_search(*inst._args(), _p(node, C.c_int32), C.c_int(node.size), C.c_int(families),
                              C.c_int(cap_keys), C.c_int(threads), _p(keys, C.c_uint64), C.byref(cost))
    return node, keys[:m].copy(), int(cost.value)


BATCH_ROUNDS = 3   # SPEC §8.1


def select_batch(inst, node, families=ALL_FAMILIES, threads=1, rounds=BATCH_ROUNDS):
    """SPEC §8.1: (round-1 per-route minimum keys m[R], selected keys round after round)."""
    node = np.ascontiguousarray(node, dtype=np.int32)
    R = node.size - inst.n
    m = np.zeros(R, np.uint64)
    keys = np.zeros(max(R, 1), np.uint64)
    c = lib().or_select_batch(*inst._args(), _p(node, C.c_int32), C.c_int(node.size), C.c_int(families),
                              C.c_int(rounds), C.c_int(threads), _p(m, C.c_uint64), _p(keys, C.c_uint64))
    return m, keys[:c].copy()


def apply_batch(inst, node, keys):
    node = np.array(node, dtype=np.int32)
    keys = np.ascontiguousarray(keys, dtype=np.uint64)
    lib().or_apply_batch(C.c_int(inst.n), _p(node, C.c_int32), C.c_int(node.size), _p(keys, C.c_uint64),
                         C.c_int(keys.size))
    return node


def local_search_batched(inst, node, families=ALL_FAMILIES, max_iters=1 << 30, threads=1, rounds=BATCH_ROUNDS):
    """SPEC §8.1.  Returns (final node[], applied keys batch after batch, batch sizes, final cost)."""
    node = np.array(node, dtype=np.int32)
    cap = 1 << 20
    keys = np.zeros(cap, np.uint64)
    sizes = np.zeros(cap, np.int32)
    iters, cost = C.c_int32(0), C.c_int64(0)
    m = lib().or_local_search_batched(*inst._args(), _p(node, C.c_int32), C.c_int(node.size), C.c_int(families),
                                      C.c_int(rounds), C.c_int(min(max_iters, cap)), C.c_int(threads), _p(keys, C.c_uint64),
                                      C.c_int(cap), _p(sizes, C.c_int32), C.c_int(cap), C.byref(iters), C.byref(cost))
    return node, keys[:m].copy(), sizes[:iters.value].copy(), int(cost.value)


# ----------------------------------------------------------------------------------------------- SPEC §9

def eta_features(distance_m, weather, traffic, weekday, hour, age):
    distance_m = np.ascontiguousarray(distance_m, dtype=np.float64)
    n = distance_m.size
    cols = [np.ascontiguousarray(a, dtype=np.uint8) for a in (weather, traffic, weekday, hour)]
    age = np.ascontiguousarray(age, dtype=np.float32)
    rows = np.zeros((n, 12), np.float32)
    lib().or_eta_features(_p(distance_m, C.c_double), *[_p(c, C.c_uint8) for c in cols], _p(age, C.c_float),
                          C.c_int64(n), _p(rows, C.c_float))
    return rows


def eta_eval(blob: bytes, rows, threads=1):
    rows = np.ascontiguousarray(rows, dtype=np.float32)
    out = np.zeros(rows.shape[0], np.float32)
    buf = np.frombuffer(blob, dtype=np.uint8)
    lib().or_eta_eval.restype = C.c_int
    rc = lib().or_eta_eval(_p(buf, C.c_uint8), C.c_int64(buf.size), _p(rows, C.c_float), C.c_int64(rows.shape[0]),
                           _p(out, C.c_float), C.c_int(threads))
    assert rc == 0, "bad model blob"
    return out
