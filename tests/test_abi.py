"""CPU: the C-ABI library loads and exports exactly what include/routest.h declares (no compute calls here)."""
import ctypes
import os
import re

import pytest

from conftest import ROOT
from routest_b200 import _native as N


def _header_functions():
    src = open(os.path.join(ROOT, "include", "routest.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(rt_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    names = _header_functions()
    assert names, "no declarations found in include/routest.h"
    handle = ctypes.CDLL(N.LIB_PATH)
    for name in names:
        assert hasattr(handle, name), f"{name} is declared in routest.h but not exported by libroutest_cuda.so"


def test_python_binding_covers_the_header():
    assert sorted(N.SIGNATURES) == _header_functions()


def test_version_and_pitch_need_no_gpu():
    lib = N.lib()
    assert lib.rt_version() >= 100
    assert lib.rt_matrix_pitch(1001) == 1008 and lib.rt_matrix_pitch(16) == 16 and lib.rt_matrix_pitch(0) == 0


def test_create_fails_loudly_without_a_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from routest_b200 import Engine, EngineError
    with pytest.raises(EngineError) as ei:
        Engine(0)
    assert ei.value.code == N.RT_ERR_NO_DEVICE and "no CPU fallback" in str(ei.value)
