"""routest_b200 — B200-native engine for the route optimizer's data-parallel hot path.

Host side in Python (the reference service's own language) over a C ABI (include/routest.h) implemented in
hand-written sm_100a CUDA (routest_b200/csrc).  There is no CPU fallback: without the built library or without a
GPU every entry point raises EngineError.
"""
from .engine import (ALL_FAMILIES, NO_MOVE, Engine, EngineError, Problem, PROFILE_SPEED_MPS, key_unpack,  # noqa: F401
                     profile_for_vehicle, time_scale_for_profile)

__all__ = ["Engine", "EngineError", "PROFILE_SPEED_MPS", "profile_for_vehicle", "time_scale_for_profile"]
