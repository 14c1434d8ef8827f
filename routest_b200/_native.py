"""Loads routest_b200/libroutest_cuda.so and declares the C ABI of include/routest.h for ctypes.

The library is built in-tree by `make -C routest_b200/csrc` (see __graft_entry__.build).  Loading fails loudly:
there is no fallback implementation.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libroutest_cuda.so")
_lib = None

RT_OK = 0
RT_ERR_INVALID_ARG = 1
RT_ERR_CUDA = 2
RT_ERR_NO_DEVICE = 3
RT_ERR_OUT_OF_MEMORY = 4
RT_ERR_INFEASIBLE_STOP = 5
RT_ERR_STATE = 6

c_engine_p = C.c_void_p
c_f64_p = C.POINTER(C.c_double)
c_i32_p = C.POINTER(C.c_int32)

# name -> (restype, argtypes); mirrors include/routest.h one to one (tests/test_abi.py checks the header against it)
SIGNATURES = {
    "rt_version": (C.c_int, []),
    "rt_create": (C.c_int, [C.c_int, C.POINTER(c_engine_p)]),
    "rt_destroy": (None, [c_engine_p]),
    "rt_last_error": (C.c_char_p, [c_engine_p]),
    "rt_engine_stream": (C.c_void_p, [c_engine_p]),
    "rt_host_alloc": (C.c_int, [c_engine_p, C.c_size_t, C.POINTER(C.c_void_p)]),
    "rt_host_free": (C.c_int, [c_engine_p, C.c_void_p]),
    "rt_matrix_pitch": (C.c_int, [C.c_int]),
    "rt_matrix_build": (C.c_int, [c_engine_p, c_f64_p, c_f64_p, C.c_int, C.c_int, C.c_int, C.c_double, C.c_double,
                                  c_i32_p, c_i32_p, c_f64_p]),
    "rt_selftest_fast_sqrt": (C.c_int, [c_engine_p, C.c_uint32, C.c_uint32, C.POINTER(C.c_uint64)]),
    "rt_unit_vectors_dev": (C.c_int, [c_engine_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
    "rt_matrix_cells_dev": (C.c_int, [c_engine_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_double,
                                      C.c_double, C.c_void_p, C.c_void_p]),
}


def build(verbose: bool = False) -> str:
    """Compiles the CUDA library in-tree for sm_100a (no GPU needed)."""
    out = None if verbose else subprocess.DEVNULL
    subprocess.check_call(["make", "-C", os.path.join(_HERE, "csrc")], stdout=out)
    return LIB_PATH


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise OSError(f"{LIB_PATH} is missing: build it with `make -C routest_b200/csrc` "
                          "(there is no fallback implementation)")
        handle = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib
