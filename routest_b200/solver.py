"""Solver front end — mirrors Flaskr/utils.py of the reference (same function names, argument meaning, defaults and
error strings) with the matrix and the trip construction done by the engine.

Reference                         here
utils.optimize_route   :10-48     optimize_route          same dict in, same dict out ({"error": ...} on failure)
utils._point_to_point  :53-82     _point_to_point         capacity / maximum_distance checks at :68-80
utils._multi_stop      :85-193    _multi_stop             :94-109 matrix -> engine, :111-139 greedy -> engine,
                                                          :141-165 directions -> hook (offline stand-in by default),
                                                          :167-193 assembly unchanged
utils._annotate_common_props :196-201

Hooks (both optional) keep the remote services pluggable exactly where the reference calls them:
  matrix_fn(points_coords, profile_type) -> {"distances": [[...]]} | {"error": str}      (utils.py:97-109)
  directions_fn(coords, profile_type)    -> GeoJSON Feature        | {"error": str}      (utils.py:55-66, 151-160)
Without hooks the matrix is the engine's great-circle matrix (SPEC.md §2-§3, kept on the GPU) and directions are the
straight-line stand-in below, so a solve needs no network at all.

Optional request fields (absent => reference behaviour): per destination `tw_start`, `tw_end` (seconds from the start
of the shift) and `service_time` (seconds); `driver_details.shift_end` (seconds, default 43200);
`driver_details.improve` runs a local search on the greedy trips: `true` = best improvement, one move per iteration
(SPEC.md §8); `"batched"` = a set of independent moves per iteration (SPEC.md §8.1; about ten times faster at 5,000
stops, ends in a different local optimum of the same neighbourhood).
"""
from __future__ import annotations

import math
import os
import threading

import numpy as np

from . import _native as N
from .engine import DEFAULT_DIST_SCALE, Engine, EngineError, profile_for_vehicle, time_scale_for_profile

ENGINE_LABEL = "backend:b200"

_engine_lock = threading.Lock()
_engine = None


def get_engine() -> Engine:
    """Process-wide engine, created lazily under a lock (handlers run on arbitrary worker threads, app.py:8)."""
    global _engine
    with _engine_lock:
        if _engine is None:
            _engine = Engine(int(os.getenv("ROUTEST_GPU", "0")))
        return _engine


def optimize_route(input_data: dict, *, engine: Engine | None = None, matrix_fn=None, directions_fn=None,
                   engine_label: str = ENGINE_LABEL):
    """Pure function: returns a Python dict (never a Flask Response) — utils.py:10-48."""
    if not input_data or not input_data.get("destination_points"):
        return {"error": "no destination points specified."}                      # utils.py:15-16

    driver_details = input_data.get("driver_details") or {}                        # utils.py:18
    vehicle_type = (driver_details.get("vehicle_type") or "car").lower().strip()   # utils.py:19
    profile_type = profile_for_vehicle(vehicle_type)                               # utils.py:22-29

    source = input_data["source_point"]                                            # utils.py:31
    destinations = input_data["destination_points"]
    hooks = dict(engine=engine, matrix_fn=matrix_fn, directions_fn=directions_fn)

    # What the reference does for a failed ORS request (utils.py:63-66, 106-109, 157-160: the exception becomes an
    # {"error": ...} dict) is done here for a failed engine call: no GPU, a coordinate out of range, a size limit.
    try:
        return _optimize(source, destinations, profile_type, driver_details, vehicle_type, engine_label, hooks)
    except EngineError as exc:
        return {"error": f"route engine failed: {exc.message}"}


def _optimize(source, destinations, profile_type, driver_details, vehicle_type, engine_label, hooks):
    if len(destinations) == 1:                                                     # utils.py:34-43
        feature = _point_to_point(source, destinations[0], profile_type, driver_details, **hooks)
        if "error" in feature:
            return feature
        p = feature.setdefault("properties", {})
        p["optimized_order"] = [0]
        p["source"] = source
        p["destinations"] = [destinations[0]]
        _annotate_common_props(feature, driver_details, vehicle_type, engine=engine_label)
        return feature

    feature = _multi_stop(source, destinations, profile_type, driver_details, **hooks)   # utils.py:45-48
    if "error" in feature:
        return feature
    _annotate_common_props(feature, driver_details, vehicle_type, engine=engine_label)
    return feature


# ---------- helpers ----------

def _point_to_point(source, destination, profile_type, driver_details, *, engine=None, matrix_fn=None,
                    directions_fn=None):
    coordinates = [[source["lon"], source["lat"]], [destination["lon"], destination["lat"]]]   # utils.py:54
    if directions_fn is None:
        directions_fn = _OfflineDirections(engine or get_engine(), profile_type)
    feature = directions_fn(coordinates, profile_type)
    if "error" in feature:
        return feature

    # Basic feasibility checks (utils.py:68-80)
    payload = destination.get("payload", 0)
    cap = driver_details.get("vehicle_capacity", 999999)
    max_dist = float(driver_details.get("maximum_distance", 9e12))
    dist_m = float(feature["properties"]["summary"]["distance"])
    errors = []
    if payload > cap:
        errors.append("payload exceeds vehicle capacity")
    if dist_m > max_dist:
        errors.append("route distance exceeds maximum_distance")
    if errors:
        return {"error": " | ".join(errors)}
    return feature


def _wants_extension(destinations, driver_details) -> bool:
    return bool(driver_details.get("improve")) or any(("tw_start" in d or "tw_end" in d or "service_time" in d)
                                                      for d in destinations)


def _multi_stop(source, destinations, profile_type, driver_details, *, engine=None, matrix_fn=None,
                directions_fn=None):
    """Capacity-aware greedy routing over a distance matrix, then per-trip directions — utils.py:85-193."""
    all_points = [source] + destinations                                           # utils.py:94
    points_coords = [[p["lon"], p["lat"]] for p in all_points]                     # utils.py:95
    n = len(all_points)

    cap = float(driver_details.get("vehicle_capacity", 9e12))                      # utils.py:112
    max_dist = float(driver_details.get("maximum_distance", 9e12))                 # utils.py:113
    demand = np.array([0.0] + [float(p.get("payload", 0)) for p in destinations])  # utils.py:126

    eng = engine
    offline = None
    if matrix_fn is not None:
        # a matrix service answered (utils.py:97-109): run the engine's greedy on exactly that matrix
        mres = matrix_fn(points_coords, profile_type)
        if "error" in mres:
            return mres
        distance_matrix = mres.get("distances")
        if not distance_matrix:
            return {"error": "ORS matrix returned no distances"}                  # utils.py:104-105
        eng = eng or get_engine()
        if _wants_extension(destinations, driver_details):
            # windows / improvement on the service's (possibly directional) matrix, integerised like the engine's own
            res = _solve_on_supplied_matrix(eng, distance_matrix, mres.get("durations"), profile_type, demand, cap,
                                            max_dist, destinations, driver_details)
            if "error" in res:
                return res
            trips = res["trips"]
        else:
            trips, infeasible = eng.greedy_trips(np.asarray(distance_matrix, dtype=np.float64), demand, cap, max_dist)
            if infeasible:
                return {"error": _infeasible_message(trips, n)}
    else:
        eng = eng or get_engine()
        res = _solve_on_device(eng, all_points, profile_type, demand, cap, max_dist, destinations, driver_details)
        if "error" in res:
            return res
        trips, offline = res["trips"], res["offline"]

    trips_indices = [[0] + list(t) + [0] for t in trips]                           # utils.py:119,135-136

    # Build directions per trip and combine (utils.py:141-165)
    if directions_fn is None:
        directions_fn = offline or _OfflineDirections(eng, profile_type)
    combined_geometry, combined_segments = [], []
    total_distance = total_duration = 0.0
    for trip in trips_indices:
        trip_coords = [[all_points[i]["lon"], all_points[i]["lat"]] for i in trip]
        feature = directions_fn(trip_coords, profile_type)
        if "error" in feature:
            return feature
        combined_geometry += feature["geometry"]["coordinates"]
        combined_segments += feature["properties"].get("segments", [])
        total_distance += float(feature["properties"]["summary"]["distance"])
        total_duration += float(feature["properties"]["summary"]["duration"])

    lons = [c[0] for c in combined_geometry]                                       # utils.py:167-169
    lats = [c[1] for c in combined_geometry]
    bbox = [min(lons), min(lats), max(lons), max(lats)]

    optimized_order = [idx - 1 for trip in trips_indices for idx in trip[1:-1]]    # utils.py:172-175

    return {                                                                       # utils.py:177-193
        "bbox": bbox,
        "type": "Feature",
        "geometry": {"type": "LineString", "coordinates": combined_geometry},
        "properties": {
            "source": source,
            "destinations": destinations,
            "optimized_order": optimized_order,
            "segments": combined_segments,
            "summary": {"distance": total_distance, "duration": total_duration, "trips": len(trips_indices)},
        },
    }


def _annotate_common_props(feature: dict, driver_details: dict, vehicle_type: str, engine: str):
    p = feature.setdefault("properties", {})                                       # utils.py:196-201
    p["vehicle_type"] = vehicle_type
    p["driver_name"] = driver_details.get("driver_name")
    p["engine"] = engine


def _infeasible_message(trips, n) -> str:
    served = {i for t in trips for i in t}
    missing = [i - 1 for i in range(1, n) if i not in served]
    return ("destination(s) %s fit no trip on their own under vehicle_capacity / maximum_distance / time windows"
            % missing[:10])


# ---------- engine-backed pieces ----------

def _solve_on_device(eng, all_points, profile_type, demand, cap, max_dist, destinations, driver_details):
    """Matrix (utils.py:94-109) and trips (utils.py:111-139) without the matrix leaving the GPU; optional time
    windows and local search (SPEC.md §4, §8)."""
    n = len(all_points)
    lat = np.array([float(p["lat"]) for p in all_points])
    lon = np.array([float(p["lon"]) for p in all_points])
    ds, ts = DEFAULT_DIST_SCALE, time_scale_for_profile(profile_type)
    extended = _wants_extension(destinations, driver_details)
    with eng.lease_problem(n) as pr:   # kept per size on the engine: no forty cudaMalloc / cudaFree per request
        pr.set_coords(lat, lon)
        pr.build_matrix(ds, ts)
        md_units = min(max_dist * ds, 9e18)
        if extended:
            stops = _stop_records(n, destinations, driver_details, demand)
            if stops is None:
                return {"error": "time windows / improve need non-negative integer payloads"}
            pr.set_stops(stops, [int(min(cap, 2 ** 28))], [int(min(md_units, 2 ** 62))])
        trips, status = pr.greedy(demand[None, :], [cap], [md_units], use_tw=extended)
        if status[0] == N.RT_ERR_INFEASIBLE_STOP:
            return {"error": _infeasible_message(trips[0], n)}
        trips = trips[0]
        if extended and driver_details.get("improve"):
            trips = _improve(pr, trips, driver_details)
        D, T = pr.get_matrix(0) if n <= 2048 else (None, None)
    return {"trips": trips, "offline": _OfflineDirections(eng, profile_type, lat=lat, lon=lon, D=D, T=T, ds=ds)}


def _improve(pr, trips, driver_details):
    pr.set_routes([trips])
    if driver_details.get("improve") == "batched" and len(trips) <= 1024:   # the batched driver's route limit
        pr.local_search(batched=True)
    else:
        pr.local_search()
    return [t for t in pr.get_routes()[0] if t]


def _stop_records(n, destinations, driver_details, demand):
    horizon = int(driver_details.get("shift_end", 43200))
    stops = np.zeros((1, n, 4), np.int32)
    stops[0, 0] = (0, horizon, 0, 0)
    for i, d in enumerate(destinations, start=1):
        dem = float(demand[i])
        if dem != int(dem) or dem < 0:
            return None
        stops[0, i] = (int(d.get("tw_start", 0)), int(d.get("tw_end", horizon)), int(d.get("service_time", 0)), int(dem))
    return stops


def _solve_on_supplied_matrix(eng, distance_matrix, durations, profile_type, demand, cap, max_dist, destinations,
                              driver_details):
    """Windows / local search on a matrix a service answered with (utils.py:97-103): metres -> integer distance units
    (x10, like the engine's own matrix), seconds from `durations` when the service sent them, else distance / profile
    speed.  The matrix may be directional."""
    D = np.asarray(distance_matrix, dtype=np.float64)
    n = D.shape[0]
    ds = DEFAULT_DIST_SCALE
    Di = np.rint(D * ds)
    Ti = np.rint(np.asarray(durations, dtype=np.float64)) if durations else np.rint(D * time_scale_for_profile(profile_type))
    if D.shape != (n, n) or Ti.shape != (n, n) or not np.isfinite(Di).all() or Di.min() < 0 or Di.max() >= 2 ** 29:
        return {"error": "matrix not usable: need a square matrix of finite, non-negative distances below 53,000 km"}
    stops = _stop_records(n, destinations, driver_details, demand)
    if stops is None:
        return {"error": "time windows / improve need non-negative integer payloads"}
    pr = eng.problem(n, 1)
    try:
        pr.set_matrix(0, Di.astype(np.int32), Ti.astype(np.int32))
        md_units = min(max_dist * ds, 9e18)
        pr.set_stops(stops, [int(min(cap, 2 ** 28))], [int(min(md_units, 2 ** 62))])
        trips, status = pr.greedy(demand[None, :], [cap], [md_units], use_tw=True)
        if status[0] == N.RT_ERR_INFEASIBLE_STOP:
            return {"error": _infeasible_message(trips[0], n)}
        trips = trips[0]
        if driver_details.get("improve"):
            trips = _improve(pr, trips, driver_details)
    finally:
        pr.close()
    return {"trips": trips}


class _OfflineDirections:
    """Stand-in for the ORS directions call (utils.py:55-62, 151-156): a straight-line LineString through the trip's
    points and one segment per leg whose distance/duration come from the engine's matrix cells (the history page of
    the frontend already draws exactly such a dashed straight-line fallback: history/[id]/page.jsx:226-239).
    Shape of what the UI reads: segments[].steps[].{instruction,type,name,distance,duration} (page.jsx:1513-1525),
    summary.{distance,duration}."""

    def __init__(self, engine, profile_type, lat=None, lon=None, D=None, T=None, ds=DEFAULT_DIST_SCALE):
        self.engine, self.profile, self.ds = engine, profile_type, ds
        self.index = {}
        self.D, self.T = D, T
        if lat is not None and D is not None:
            self.index = {(float(lo), float(la)): i for i, (la, lo) in enumerate(zip(lat, lon))}

    def _legs(self, coords):
        idx = [self.index.get((float(c[0]), float(c[1]))) for c in coords]
        if self.D is not None and all(i is not None for i in idx):
            return [(self.D[a, b] / self.ds, float(self.T[a, b])) for a, b in zip(idx[:-1], idx[1:])]
        # points the cached matrix does not know: one small matrix build for this trip
        lat = np.array([c[1] for c in coords], dtype=np.float64)
        lon = np.array([c[0] for c in coords], dtype=np.float64)
        m = self.engine.matrix_build(lat, lon, dist_scale=self.ds, time_scale=time_scale_for_profile(self.profile))
        k = range(len(coords) - 1)
        return [(m["dist"][i, i + 1] / self.ds, float(m["time"][i, i + 1])) for i in k]

    def __call__(self, coords, profile_type):
        legs = self._legs(coords)
        segments = []
        for k, (dist, dur) in enumerate(legs):
            last = k == len(legs) - 1
            steps = [{"distance": dist, "duration": dur, "type": 11, "instruction": "Head straight to the next stop",
                      "name": "-", "way_points": [k, k + 1]},
                     {"distance": 0.0, "duration": 0.0, "type": 10,
                      "instruction": "Arrive at the depot" if last else "Arrive at the stop", "name": "-",
                      "way_points": [k + 1, k + 1]}]
            segments.append({"distance": dist, "duration": dur, "steps": steps})
        total_d = math.fsum(s["distance"] for s in segments)
        total_t = math.fsum(s["duration"] for s in segments)
        lons = [c[0] for c in coords]
        lats = [c[1] for c in coords]
        return {"type": "Feature", "bbox": [min(lons), min(lats), max(lons), max(lats)],
                "geometry": {"type": "LineString", "coordinates": [list(c) for c in coords]},
                "properties": {"segments": segments, "summary": {"distance": total_d, "duration": total_t},
                               "way_points": [0, len(coords) - 1]}}
