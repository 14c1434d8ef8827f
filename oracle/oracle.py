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
    src = os.path.join(_HERE, "routest_oracle.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
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
