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
c_i64_p = C.POINTER(C.c_int64)
c_u64_p = C.POINTER(C.c_uint64)
c_u8_p = C.POINTER(C.c_uint8)
c_problem_p = C.c_void_p

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
    "rt_matrix_static_max_nodes": (C.c_int, []),
    "rt_matrix_cells_static_dev": (C.c_int, [c_engine_p, C.c_void_p, C.c_void_p, C.c_int, C.c_double, C.c_double,
                                             C.c_void_p, C.c_void_p]),
    "rt_greedy_trips": (C.c_int, [c_engine_p, c_f64_p, C.c_int, c_f64_p, C.c_double, C.c_double, c_i32_p, c_i32_p,
                                  c_i32_p]),
    "rt_problem_create": (C.c_int, [c_engine_p, C.c_int, C.c_int, C.POINTER(c_problem_p)]),
    "rt_problem_destroy": (None, [c_problem_p]),
    "rt_problem_position_stride": (C.c_int, [c_problem_p]),
    "rt_problem_set_coords": (C.c_int, [c_problem_p, c_f64_p, c_f64_p]),
    "rt_problem_set_stops": (C.c_int, [c_problem_p, c_i32_p, c_i32_p, c_i64_p]),
    "rt_problem_build_matrix": (C.c_int, [c_problem_p, C.c_double, C.c_double]),
    "rt_problem_set_matrix": (C.c_int, [c_problem_p, C.c_int, c_i32_p, c_i32_p]),
    "rt_problem_get_matrix": (C.c_int, [c_problem_p, C.c_int, c_i32_p, c_i32_p]),
    "rt_problem_greedy": (C.c_int, [c_problem_p, c_f64_p, c_f64_p, c_f64_p, C.c_int, c_i32_p, c_i32_p, c_i32_p,
                                    c_i32_p]),
    "rt_problem_set_routes": (C.c_int, [c_problem_p, c_i32_p, c_i32_p, c_i32_p]),
    "rt_problem_get_routes": (C.c_int, [c_problem_p, c_i32_p, c_i32_p, c_i32_p]),
    "rt_problem_get_state": (C.c_int, [c_problem_p, C.c_int, c_i32_p, c_i32_p, c_i32_p, c_i32_p, c_i32_p, c_i64_p,
                                       c_i64_p, c_i64_p, c_i32_p]),
    "rt_problem_get_position_matrix": (C.c_int, [c_problem_p, C.c_int, c_i32_p, c_i32_p]),
    "rt_problem_eval_moves": (C.c_int, [c_problem_p, C.c_int, C.c_int, C.c_int, c_u64_p]),
    "rt_problem_eval_dump": (C.c_int, [c_problem_p, C.c_int, C.c_int, c_i32_p, c_u8_p]),
    "rt_problem_apply_moves": (C.c_int, [c_problem_p, c_u64_p]),
    "rt_problem_local_search": (C.c_int, [c_problem_p, C.c_int, C.c_int, c_i32_p, c_i64_p, c_u64_p, C.c_int]),
    "rt_problem_plan": (C.c_int, [c_problem_p, c_f64_p, c_f64_p, c_i32_p, c_i32_p, c_i64_p, c_i32_p, c_i32_p,
                                  C.c_double, C.c_double, C.c_int, c_u64_p]),
    "rt_problem_plan_begin": (C.c_int, [c_problem_p, c_f64_p, c_f64_p, c_i32_p, c_i32_p, c_i64_p, c_i32_p, c_i32_p,
                                        C.c_double, C.c_double, C.c_int]),
    "rt_problem_plan_end": (C.c_int, [c_problem_p, c_u64_p]),
    "rt_problem_plan_upload_dev": (C.c_int, [c_problem_p, c_f64_p, c_f64_p, c_i32_p, c_i32_p, c_i64_p, c_i32_p, c_i32_p,
                                             C.c_double, C.c_double, C.c_int, C.c_void_p, C.c_void_p]),
    "rt_problem_build_matrix_dev": (C.c_int, [c_problem_p, C.c_double, C.c_double, C.c_void_p]),
    "rt_problem_refresh_dev": (C.c_int, [c_problem_p, C.c_int, C.c_void_p]),
    "rt_problem_plan_dev": (C.c_int, [c_problem_p, C.c_double, C.c_double, C.c_int, C.c_void_p, C.c_void_p]),
    "rt_problem_set_shard": (C.c_int, [c_problem_p, C.c_int, C.c_int]),
    "rt_problem_local_search_full": (C.c_int, [c_problem_p, C.c_int, C.c_int, c_i32_p, c_i64_p, c_u64_p, C.c_int]),
    "rt_problem_local_search_batched": (C.c_int, [c_problem_p, C.c_int, C.c_int, c_i32_p, c_i64_p, c_u64_p, C.c_int]),
    "rt_problem_eval_moves_dev": (C.c_int, [c_problem_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
    "rt_problem_apply_moves_dev": (C.c_int, [c_problem_p, C.c_void_p, C.c_int, C.c_void_p]),
    "rt_problem_table_entries": (C.c_int64, [c_problem_p]),
    "rt_problem_eval_table_dev": (C.c_int, [c_problem_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
    "rt_problem_apply_batch_dev": (C.c_int, [c_problem_p, C.c_void_p, C.c_void_p]),
    "rt_problem_exchange_handle_bytes": (C.c_int, []),
    "rt_problem_exchange_export": (C.c_int, [c_problem_p, C.c_void_p]),
    "rt_problem_exchange_connect": (C.c_int, [c_problem_p, C.c_void_p, C.c_int, C.c_int]),
    "rt_problem_eval_moves_xchg_dev": (C.c_int, [c_problem_p, C.c_int, C.c_void_p]),
    "rt_problem_apply_moves_xchg_dev": (C.c_int, [c_problem_p, C.c_int, C.c_void_p]),
    "rt_problem_collect_keys_xchg_dev": (C.c_int, [c_problem_p, C.c_void_p, C.c_void_p]),
    "rt_problem_table_exchange_export": (C.c_int, [c_problem_p, C.c_int, C.c_void_p]),
    "rt_problem_table_exchange_connect": (C.c_int, [c_problem_p, C.c_void_p, C.c_int, C.c_int]),
    "rt_problem_eval_table_xchg_dev": (C.c_int, [c_problem_p, C.c_int, C.c_void_p, C.c_void_p]),
    "rt_problem_apply_batch_xchg_dev": (C.c_int, [c_problem_p, C.c_void_p, C.c_void_p]),
    "rt_problem_search_status": (C.c_int, [c_problem_p, c_i32_p, c_i32_p, c_i64_p]),
    "rt_eta_load_model": (C.c_int, [c_engine_p, C.c_void_p, C.c_size_t]),
    "rt_eta_model_info": (C.c_int, [c_engine_p, c_i32_p, c_i32_p]),
    "rt_eta_score": (C.c_int, [c_engine_p, c_f64_p, c_u8_p, c_u8_p, c_u8_p, c_u8_p, C.POINTER(C.c_float), C.c_int64,
                               C.POINTER(C.c_float)]),
    "rt_eta_features": (C.c_int, [c_engine_p, c_f64_p, c_u8_p, c_u8_p, c_u8_p, c_u8_p, C.POINTER(C.c_float), C.c_int64,
                                  C.POINTER(C.c_float)]),
    "rt_eta_score_dev": (C.c_int, [c_engine_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                   C.c_int64, C.c_void_p, C.c_void_p]),
    "rt_selftest_fast_sqrt": (C.c_int, [c_engine_p, C.c_uint32, C.c_uint32, C.POINTER(C.c_uint64)]),
    "rt_selftest_pipe_peak": (C.c_int, [c_engine_p, C.c_int, C.c_int, c_f64_p, c_f64_p]),
    "rt_debug_set": (C.c_int, [C.c_char_p, C.c_int]),
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
