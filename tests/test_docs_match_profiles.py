"""The headline figures quoted in README.md / DESIGN.md / BASELINE.md are the ones in the committed bench lines
(profiles/r2_bench_n*.json): a number changed in one place and not the other fails here."""
import json
import os

import pytest

from conftest import ROOT


def _line(n):
    path = os.path.join(ROOT, "profiles", f"r2_bench_n{n}.json")
    for raw in open(path):
        if raw.startswith("{"):
            return json.loads(raw)
    raise AssertionError(f"no JSON line in {path}")


@pytest.mark.parametrize("n", [1, 2, 4, 8])
def test_pass_times_quoted_in_the_docs(n):
    d = _line(n)
    # (BASELINE.json's metric names the whole sweep, "... at 1/2/4/8 B200 vs host CPU"; a line carries the quantity)
    assert d["n_gpus"] == n and json.load(open(os.path.join(ROOT, "BASELINE.json")))["metric"].startswith(d["metric"])
    ms = f"{d['ms_per_step']:.4f}"
    for doc in ("README.md", "DESIGN.md", "BASELINE.md"):
        assert ms in open(os.path.join(ROOT, doc)).read(), f"{doc} does not quote {ms} ms per pass at N = {n}"


def test_bench_lines_carry_the_contract_keys():
    for n in (1, 2, 4, 8):
        d = _line(n)
        for key in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
                    "vs_baseline", "dtype", "data", "config", "roofline", "clocks", "e2e", "gpu_launches"):
            assert key in d, (n, key)
        assert d["scaling"] == "strong" and d["gpu_launches"] > 0 and d["vs_baseline"] is None
        r = d["roofline"]
        assert r["bound"] in ("issue", "alu") and 0.0 < r["frac"] < 1.2
        assert r["matrix_kernel"]["bound"] == "hbm" and 0.0 < r["matrix_kernel"]["frac"] < 1.2
        e = d["e2e"]
        assert e["h2d_bytes_per_step"] > 0 and e["d2h_bytes_per_step"] > 0 and e["value"] < d["value"]
        assert not set(d["clocks"]["reasons"]) & {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"}
    one = _line(1)
    assert one["cpu_baseline"]["kind"] == "port" and one["cpu_baseline"]["cores"] >= 1
    assert one["extra"]["batch_64x1000"]["keys_equal_oracle"] is True
