"""Engine — Python face of the C ABI (include/routest.h).

One Engine per GPU and per process.  Methods without a `_dev` suffix take and return numpy arrays (host memory) and
do their own copies; `_dev` methods take raw device pointers (e.g. `torch.Tensor.data_ptr()`) and a CUDA stream
handle and never synchronise — torch supplies device memory, streams and NCCL, nothing else.
"""
from __future__ import annotations

import ctypes as C

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

    # -- lifetime -----------------------------------------------------------------------------------------------
    def close(self):
        if getattr(self, "_h", None):
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

    def unit_vectors_dev(self, lat_ptr: int, lon_ptr: int, n: int, batch: int, xyz_ptr: int, stream: int = 0):
        self._check(self._lib.rt_unit_vectors_dev(self._h, lat_ptr, lon_ptr, int(n), int(batch), xyz_ptr, stream or None))

    def matrix_cells_dev(self, xyz_ptr: int, n: int, batch: int, cells_ptr: int, *, rows=None,
                         dist_scale: float = DEFAULT_DIST_SCALE, time_scale: float | None = None, stream: int = 0):
        r0, r1 = rows if rows is not None else (0, n)
        if time_scale is None:
            time_scale = time_scale_for_profile("driving-car")
        self._check(self._lib.rt_matrix_cells_dev(self._h, xyz_ptr, int(n), int(batch), int(r0), int(r1),
                                                  float(dist_scale), float(time_scale), cells_ptr, stream or None))
