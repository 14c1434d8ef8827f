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


def test_every_debug_knob_is_documented_in_the_header():
    """rt_debug_set takes its knobs by name; every name the library accepts (csrc/moves.cu) is described in the
    comment above its declaration in include/routest.h, and the header describes none that does not exist."""
    import re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    src = open(os.path.join(root, "routest_b200", "csrc", "moves.cu")).read()
    body = src[src.index('extern "C" int rt_debug_set'):]
    accepted = set(re.findall(r'strcmp\(name, "([a-z_]+)"\)', body))
    assert {"strip_rows", "fused_eval", "persistent_search", "apply_prep"} <= accepted
    hdr = open(os.path.join(root, "include", "routest.h")).read()
    doc = hdr[hdr.index("/* Test and measurement knobs"):hdr.index("int rt_debug_set")]
    documented = set(re.findall(r'"([a-z_]+)"', doc))
    assert accepted == documented, (sorted(accepted - documented), sorted(documented - accepted))
