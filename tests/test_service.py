"""Handler mirrors (routest_b200/service.py) against what routes.py:29-50, 89-127, 365-383 do with the same bodies."""
import datetime as dt
import json

import pytest

from conftest import load_golden
from oracle.ref_loader import stub_directions
from routest_b200 import service as S

CASES = {c["name"]: c for c in load_golden("optimize_route_cases.json")}


def _hooks(case):
    return dict(matrix_fn=lambda coords, profile: {"distances": case["matrix"]},
                directions_fn=lambda coords, profile: stub_directions(coords), engine_label="backend:ors")


def test_request_route_returns_errors_with_200():
    body, status = S.request_route({}, **_hooks(CASES["empty_body"]))
    assert status == 200 and body == {"error": "no destination points specified."}   # routes.py:45-50


def test_optimize_route_endpoint_errors_are_400():
    body, status = S.optimize_route_endpoint(None, **_hooks(CASES["empty_body"]))
    assert status == 400 and body == {"error": "no destination points specified."}   # routes.py:93-94


def test_optimize_route_endpoint_attaches_ml_eta():
    case = CASES["single_stop_ok"]
    seen = {}

    def predict(**kw):
        seen.update(kw)
        return 12.5, "2025-01-01T00:12:30"

    payload = json.loads(json.dumps(case["body"]))
    payload.update(use_ml_eta=True, context={"traffic": "Medium"})
    payload["driver_details"]["driver_age"] = 44
    now = dt.datetime(2025, 1, 1)
    body, status = S.optimize_route_endpoint(payload, now=now, predict=predict, **_hooks(case))
    assert status == 200
    assert body["properties"]["eta_minutes_ml"] == 12.5 and body["properties"]["eta_completion_time_ml"] == "2025-01-01T00:12:30"
    assert seen == dict(weather="Sunny", traffic="Medium", distance_m=case["result"]["properties"]["summary"]["distance"],
                        pickup_time=now, driver_age=44.0)                            # routes.py:99-112
    # model unavailable: fields simply absent (routes.py:114-116)
    body, _ = S.optimize_route_endpoint(payload, now=now, predict=lambda **kw: (None, None), **_hooks(case))
    assert "eta_minutes_ml" not in body["properties"]
    # not requested: the reference result unchanged
    plain, _ = S.optimize_route_endpoint(json.loads(json.dumps(case["body"])), **_hooks(case))
    assert plain == case["result"]


def test_predict_eta_endpoint_defaults_and_503():
    seen = {}

    def predict(**kw):
        seen.update(kw)
        return 7.0, "x"

    now = dt.datetime(2025, 2, 3, 4, 5, 6)
    body, status = S.predict_eta_endpoint({"summary": {"distance": 2500}}, now=now, predict=predict)
    assert (body, status) == ({"eta_minutes_ml": 7.0, "eta_completion_time_ml": "x"}, 200)
    assert seen == dict(weather="Sunny", traffic="Low", distance_m=2500.0, pickup_time=now.isoformat(), driver_age=30.0)
    assert S.predict_eta_endpoint(None, predict=lambda **kw: (None, None)) == ({"error": "model unavailable"}, 503)


@pytest.mark.gpu
def test_endpoints_on_the_engine(engine):
    """Offline path end to end: engine matrix + greedy + stand-in directions + engine ETA."""
    from test_eta import synth_model
    from routest_b200 import ml as M
    _, blob = synth_model(seed=6)
    M.use_engine(engine, blob)
    payload = json.loads(json.dumps(CASES["eight_stops_float_matrix"]["body"]))
    payload.update(use_ml_eta=True, context={"weather": "Stormy", "traffic": "Jam"})
    payload["driver_details"].pop("maximum_distance")   # the golden body's limit was sized for its stub matrix
    body, status = S.optimize_route_endpoint(payload, engine=engine)
    assert status == 200 and body["properties"]["engine"] == "backend:b200"
    assert body["properties"]["eta_minutes_ml"] > 0 and "T" in body["properties"]["eta_completion_time_ml"]
    eta, st = S.predict_eta_endpoint({"summary": {"distance": body["properties"]["summary"]["distance"]},
                                      "pickup_time": "2025-08-12T14:30:00", "weather": "Stormy", "traffic": "Jam"})
    assert st == 200 and eta["eta_minutes_ml"] > 0
    M._state.update(engine=None, loaded=None)
