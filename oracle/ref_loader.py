"""Imports the reference's own solver/ETA modules BY FILE PATH with the network stubbed out.

Works only where /root/reference exists (the build container).  It is used by tests/golden/make_golden.py to
freeze golden vectors and by tests that are skipped when the reference is absent; nothing that runs on the GPU
box reads /root/reference.  TEST INFRASTRUCTURE ONLY.

`import Flaskr` would execute Flaskr/__init__.py, which needs flask/flask_cors/flask_sse (absent here), so the two
modules are loaded straight from their files.
"""
from __future__ import annotations

import importlib.util
import os

REF_ROOT = "/root/reference/backend/route_optimizer_twx2/Flaskr"


def available() -> bool:
    return os.path.isfile(os.path.join(REF_ROOT, "utils.py"))


def _load(name: str):
    spec = importlib.util.spec_from_file_location(f"_routest_ref_{name}", os.path.join(REF_ROOT, f"{name}.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


class _Resp:
    def __init__(self, payload):
        self._payload = payload

    def raise_for_status(self):
        return None

    def json(self):
        return self._payload


def stub_directions(coords):
    """Deterministic stand-in for an ORS directions answer: straight segments, Manhattan degrees x 1e5 as metres,
    10 m/s.  Used on both sides of the wrapper-parity tests (reference via StubORS, solver via directions_fn)."""
    segs, td, tt = [], 0.0, 0.0
    for a, b in zip(coords[:-1], coords[1:]):
        d = round((abs(a[0] - b[0]) + abs(a[1] - b[1])) * 1e5, 1)
        segs.append({"distance": d, "duration": d / 10.0,
                     "steps": [{"distance": d, "duration": d / 10.0, "type": 11, "instruction": "Head on", "name": "-"}]})
        td += d
        tt += d / 10.0
    return {"type": "Feature", "geometry": {"type": "LineString", "coordinates": [list(c) for c in coords]},
            "properties": {"segments": segs, "summary": {"distance": td, "duration": tt}}}


class StubORS:
    """Stands in for requests.post: serves a caller-supplied matrix for /matrix/ URLs and a straight-line
    Feature for /directions/ URLs, and records the calls."""

    def __init__(self, matrix):
        self.matrix = matrix
        self.calls = []

    def __call__(self, url, json=None, headers=None, timeout=None):
        self.calls.append((url, json))
        if "/matrix/" in url:
            return _Resp({"distances": self.matrix})
        return _Resp({"features": [stub_directions(json["coordinates"])]})


def load_utils(matrix):
    """Returns (module, stub) — the reference utils.py with requests.post replaced by a StubORS(matrix)."""
    mod = _load("utils")
    stub = StubORS(matrix)

    class _Requests:
        RequestException = mod.requests.RequestException
        post = staticmethod(stub)

    mod.requests = _Requests
    return mod, stub


def load_ml():
    return _load("ml")


def reference_trips(matrix, source, destinations, driver_details, profile="driving-car"):
    """Runs the reference's _multi_stop (utils.py:85-193) on `matrix` and returns its Feature plus the
    trips recovered from the recorded /directions/ calls (each call's coordinates are one trip)."""
    mod, stub = load_utils(matrix)
    feat = mod._multi_stop(source, destinations, profile, driver_details)
    return feat, stub
