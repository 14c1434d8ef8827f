"""Framework-free mirrors of the three handlers on the hot path (Flaskr/routes.py): each takes the parsed JSON body and
returns `(body_dict, http_status)` exactly as the Flask view would jsonify it.  Persistence (routes.py:119-125),
SSE and health checks are out of scope and are not reproduced; plugging these into the Flask
blueprint is `return jsonify(body), status`.
"""
from __future__ import annotations

import datetime as dt

from . import ml as _ml
from . import solver as _solver


def request_route(data, **solver_kw):
    """POST /api/request_route (routes.py:29-50): an {"error": ...} dict is truthy, so it goes out with HTTP 200
    there, unlike /optimize_route."""
    response = _solver.optimize_route(data, **solver_kw)
    if not response:
        return {"error": "no response acquired from the optimizer."}, 400
    return response, 200


def optimize_route_endpoint(payload, *, now=None, predict=None, **solver_kw):
    """POST /api/optimize_route (routes.py:89-127) without the persistence step."""
    payload = payload or {}                                                   # routes.py:91
    result = _solver.optimize_route(payload, **solver_kw)
    if isinstance(result, dict) and result.get("error"):                      # routes.py:93-94
        return result, 400
    if payload.get("use_ml_eta"):                                             # routes.py:97-116
        props = result.setdefault("properties", {}) or {}
        summary = props.get("summary", {}) or {}
        distance_m = float(summary.get("distance") or 0)
        ctx = payload.get("context") or {}
        weather = ctx.get("weather", "Sunny")
        traffic = ctx.get("traffic", "Low")
        driver_age = float((payload.get("driver_details") or {}).get("driver_age", 30))
        eta_min, eta_iso = (predict or _ml.predict_eta_minutes)(
            weather=weather, traffic=traffic, distance_m=distance_m, pickup_time=now or dt.datetime.now(),
            driver_age=driver_age)
        if eta_min is not None:
            props["eta_minutes_ml"] = eta_min
            props["eta_completion_time_ml"] = eta_iso
    return result, 200


def predict_eta_endpoint(body, *, now=None, predict=None):
    """POST /api/predict_eta (routes.py:365-383)."""
    body = body or {}
    summary = body.get("summary") or {}
    pickup = body.get("pickup_time") or (now or dt.datetime.now()).isoformat()
    driver_age = float(body.get("driver_age", 30))
    weather = body.get("weather", "Sunny")
    traffic = body.get("traffic", "Low")
    eta_min, eta_iso = (predict or _ml.predict_eta_minutes)(
        weather=weather, traffic=traffic, distance_m=float(summary.get("distance") or 0), pickup_time=pickup,
        driver_age=driver_age)
    if eta_min is None:
        return {"error": "model unavailable"}, 503
    return {"eta_minutes_ml": eta_min, "eta_completion_time_ml": eta_iso}, 200
