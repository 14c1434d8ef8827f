import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")

BOX = (14.4043, 14.6656, 120.9634, 121.0891)  # Metro Manila, SURVEY.md §8(d)
DEPOT = (14.5836, 121.0409)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


def synth_instance(seed: int, n_stops: int, vehicles: int):
    """SURVEY.md §8(d) synthetic instance (same recipe as tests/golden/make_golden.py)."""
    rng = np.random.default_rng(seed)
    lat = np.concatenate([[DEPOT[0]], rng.uniform(BOX[0], BOX[1], n_stops)])
    lon = np.concatenate([[DEPOT[1]], rng.uniform(BOX[2], BOX[3], n_stops)])
    demand = np.concatenate([[0], rng.integers(1, 10, n_stops)]).astype(np.int64)
    cap = int(np.ceil(1.05 * demand.sum() / vehicles))
    return lat, lon, demand, cap


def load_golden(name):
    with open(os.path.join(GOLDEN, name)) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def engine():
    from routest_b200 import Engine
    eng = Engine(0)  # raises if the library or the GPU is missing: GPU tests must not pass on a fallback
    yield eng
    eng.close()
