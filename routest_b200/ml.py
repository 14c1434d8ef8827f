"""ETA model wrapper — mirrors Flaskr/ml.py of the reference on top of the engine's batched scorer.

`predict_eta_minutes` keeps the reference's keyword-only signature, defaults and return convention
(ml.py:23-58): `(minutes, iso_completion_time)` or `(None, None)` when no model is available.  The model is not a
pickle (ml.py:6-21) but an ensemble in the engine's tree format (SPEC.md §9.2); `export_sklearn` and
`export_xgboost_json` (XGBoost `.json` and `.ubj` dumps) write that format.  `ETA_MODEL_PATH` names the file, as in the reference (ml.py:7).
"""
from __future__ import annotations

import datetime as dt
import json
import os
import struct
import threading

import numpy as np

FEATURES = ["weather_Cloudy", "weather_Stormy", "weather_Sunny", "weather_Windy", "traffic_High", "traffic_Jam",
            "traffic_Low", "traffic_Medium", "weekday_ordered", "hour_ordered", "distance_km", "driver_age"]  # ml.py:35-48
WEATHER = {"Cloudy": 0, "Stormy": 1, "Sunny": 2, "Windy": 3}   # ml.py:36-39 (case-sensitive string equality)
TRAFFIC = {"High": 0, "Jam": 1, "Low": 2, "Medium": 3}         # ml.py:40-43
NONE_CODE = 255

_lock = threading.Lock()
_state = {"engine": None, "loaded": None}   # loaded: True / "ERROR:..." / None  (ml.py:4,11-21)


def encode_weather(w) -> int:
    return WEATHER.get(w, NONE_CODE) if isinstance(w, str) else NONE_CODE


def encode_traffic(t) -> int:
    return TRAFFIC.get(t, NONE_CODE) if isinstance(t, str) else NONE_CODE


def _model_path():
    return os.getenv("ETA_MODEL_PATH") or os.path.join(os.path.dirname(__file__), "..", "eta_model.rte")  # ml.py:6-9


def use_engine(engine, blob: bytes | None = None):
    """Binds the module-level predictor to an engine (and optionally loads a model blob right away)."""
    with _lock:
        _state["engine"] = engine
        _state["loaded"] = None
        if blob is not None:
            engine.eta_load_model(blob)
            _state["loaded"] = True


def _load_model():
    """Lazy, locked (the reference's lazy load is unlocked under threaded Flask); a failure is
    remembered as a string so later calls return (None, None) without retrying (ml.py:19-21)."""
    with _lock:
        if _state["loaded"] is not None:
            return _state["loaded"]
        try:
            if _state["engine"] is None:
                from .engine import Engine
                _state["engine"] = Engine(int(os.getenv("ROUTEST_GPU", "0")))
            with open(_model_path(), "rb") as f:
                _state["engine"].eta_load_model(f.read())
            _state["loaded"] = True
        except Exception as e:  # noqa: BLE001 - mirrors ml.py:19
            _state["loaded"] = f"ERROR:{e}"
        return _state["loaded"]


def predict_eta_minutes(*, weather: str, traffic: str, distance_m: float, pickup_time, driver_age: float = 30.0):
    if _load_model() is not True:                       # ml.py:24-26
        return None, None
    if isinstance(pickup_time, str):                    # ml.py:28-33
        pickup_dt = dt.datetime.fromisoformat(pickup_time)
    elif isinstance(pickup_time, dt.datetime):
        pickup_dt = pickup_time
    else:
        pickup_dt = dt.datetime.now()
    try:
        minutes = _state["engine"].eta_score(
            [float(distance_m or 0)], [encode_weather(weather)], [encode_traffic(traffic)], [pickup_dt.weekday()],
            [pickup_dt.hour], [float(driver_age or 30.0)])
        eta_minutes = float(minutes[0])                 # ml.py:53
    except Exception:  # noqa: BLE001 - ml.py:54-55
        return None, None
    eta_ts = (pickup_dt + dt.timedelta(minutes=eta_minutes)).isoformat()   # ml.py:57
    return eta_minutes, eta_ts


def predict_eta_batch(engine, *, weather, traffic, distance_m, pickup_times, driver_age):
    """The batch form the reference lacks: sequences of the same five inputs -> float32 minutes."""
    pts = [dt.datetime.fromisoformat(p) if isinstance(p, str) else p for p in pickup_times]
    return engine.eta_score([float(d or 0) for d in distance_m], [encode_weather(w) for w in weather],
                            [encode_traffic(t) for t in traffic], [p.weekday() for p in pts], [p.hour for p in pts],
                            [float(a or 30.0) for a in driver_age])


# ------------------------------------------------------------------------------------------------ model format

def pack_model(trees, base_score: float, n_features: int = 12) -> bytes:
    """trees: list of lists of (value, left, feature, flags) with children adjacent (SPEC.md §9.2)."""
    offs, nodes = [0], bytearray()
    for tr in trees:
        for value, left, feature, flags in tr:
            nodes += struct.pack("<fHBB", float(np.float32(value)), int(left), int(feature), int(flags))
        offs.append(offs[-1] + len(tr))
    return b"RTE1" + struct.pack("<IIf", len(trees), n_features, float(np.float32(base_score))) + \
        struct.pack(f"<{len(offs)}I", *offs) + bytes(nodes)


def _relabel(children_left, children_right, emit):
    """Breadth-first renumbering that makes siblings adjacent; emit(old_index, new_left_or_0) -> node tuple."""
    order, new_id, queue = [], {0: 0}, [0]
    while queue:
        k = queue.pop(0)
        order.append(k)
        l, r = children_left[k], children_right[k]
        if l >= 0:
            new_id[l] = len(new_id)
            new_id[r] = len(new_id)
            queue += [l, r]
    return [emit(k, new_id[children_left[k]] if children_left[k] >= 0 else 0) for k in order]


def _f32_above(t: float) -> np.float32:
    """Smallest float32 strictly greater than t: x <= t (float64 threshold, float32 x)  <=>  x < _f32_above(t)."""
    f = np.float32(t)
    if float(f) > t:
        return f
    return np.nextafter(f, np.float32(np.inf))


def export_sklearn(model) -> bytes:
    """scikit-learn GradientBoostingRegressor -> RTE1.  sklearn goes left on `x <= threshold` with float32 x;
    the engine goes left on `x < value`, so each threshold becomes the next float32 above it."""
    lr = float(model.learning_rate)
    init = model.init_
    base = float(np.ravel(init.constant_)[0]) if hasattr(init, "constant_") else float(np.ravel(init.predict(np.zeros((1, model.n_features_in_))))[0])
    trees = []
    for est in model.estimators_.ravel():
        t = est.tree_

        def emit(k, new_left, t=t):
            if t.children_left[k] < 0:
                return (lr * float(t.value[k].ravel()[0]), 0, 0, 1)
            return (_f32_above(float(t.threshold[k])), new_left, int(t.feature[k]), 2)
        trees.append(_relabel(t.children_left, t.children_right, emit))
    return pack_model(trees, base, model.n_features_in_)


_IDENTITY_OBJECTIVES = {"reg:squarederror", "reg:linear", "reg:absoluteerror", "reg:pseudohubererror",
                        "reg:quantileerror", "reg:squaredlogerror"}   # prediction = base_score + sum of leaves


def read_ubjson(buf: bytes):
    """Minimal Universal Binary JSON reader (the container XGBoost's `save_model('*.ubj')` writes): big-endian
    scalars, strings, arrays and objects, including the strongly typed `[$type#count` containers it uses for the
    per-tree arrays.  Returns the same dict/list structure `json.load` gives for the `.json` dump."""
    pos = 0
    scalar = {b"i": ">b", b"U": ">B", b"I": ">h", b"l": ">i", b"L": ">q", b"d": ">f", b"D": ">d"}

    def take(n):
        nonlocal pos
        if pos + n > len(buf):
            raise ValueError("UBJSON: truncated input")
        out = buf[pos:pos + n]
        pos += n
        return out

    def value(marker=None):
        m = marker or take(1)
        while m == b"N":
            m = take(1)
        if m in scalar:
            fmt = scalar[m]
            return struct.unpack(fmt, take(struct.calcsize(fmt)))[0]
        if m == b"Z":
            return None
        if m == b"T":
            return True
        if m == b"F":
            return False
        if m == b"C":
            return take(1).decode("latin-1")
        if m in (b"S", b"H"):
            return take(length()).decode("utf-8")
        if m == b"[":
            typ, count = header()
            if typ in scalar and count is not None:      # typed numeric array: one bulk unpack
                fmt = scalar[typ]
                return list(struct.unpack(f">{count}{fmt[1]}", take(count * struct.calcsize(fmt))))
            out = []
            while count is None or len(out) < count:
                if count is None:
                    nxt = take(1)
                    if nxt == b"]":
                        break
                    out.append(value(nxt))
                else:
                    out.append(value(typ))
            return out
        if m == b"{":
            typ, count = header()
            out = {}
            while count is None or len(out) < count:
                if count is None:
                    nxt = take(1)
                    if nxt == b"}":
                        break
                    key = take(length(nxt)).decode("utf-8")
                else:
                    key = take(length()).decode("utf-8")
                out[key] = value(typ)
            return out
        raise ValueError(f"UBJSON: unknown marker {m!r} at byte {pos - 1}")

    def length(marker=None):
        n = value(marker)
        if not isinstance(n, int) or isinstance(n, bool) or n < 0:
            raise ValueError("UBJSON: bad length")
        return n

    def header():
        nonlocal pos
        typ = count = None
        if buf[pos:pos + 1] == b"$":
            pos += 1
            typ = take(1)
            if buf[pos:pos + 1] != b"#":
                raise ValueError("UBJSON: '$' without '#'")
        if buf[pos:pos + 1] == b"#":
            pos += 1
            count = length()
        return typ, count

    doc = value()
    return doc


def load_xgboost_dump(path: str) -> dict:
    """`.json` or `.ubj` model file -> the document `export_xgboost_json` takes (XGBoost picks the container by
    file extension; so does this, falling back to sniffing: a JSON object's first key starts with a quote)."""
    with open(path, "rb") as f:
        raw = f.read()
    if path.endswith(".ubj") or (not path.endswith(".json") and raw[:1] == b"{" and raw[1:2] not in (b'"', b" ", b"\n", b"}")):
        return read_ubjson(raw)
    return json.loads(raw.decode("utf-8"))


def _xgb_scalar(v) -> float:
    """XGBoost writes model parameters as strings: "5E-1", and since 3.x "[5E-1]" (one value per target)."""
    if isinstance(v, (list, tuple)):
        if len(v) != 1:
            raise ValueError("multi-target XGBoost models are not supported")
        v = v[0]
    if isinstance(v, str):
        v = v.strip()
        if v.startswith("["):
            parts = [x for x in v.strip("[]").split(",") if x.strip()]
            if len(parts) != 1:
                raise ValueError("multi-target XGBoost models are not supported")
            v = parts[0]
    return float(v)


def export_xgboost_json(path_or_dict) -> bytes:
    """XGBoost `save_model('*.json' | '*.ubj')` dump (gbtree or dart, identity-link regression objective) -> RTE1.
    XGBoost already goes left on `x < split_condition` with float32 x and routes missing values by `default_left`.
    What the engine's format cannot express is refused with ValueError rather than exported wrongly: categorical
    splits, multi-target trees, objectives whose prediction is a transform of the margin (logistic, gamma, ...)."""
    doc = path_or_dict if isinstance(path_or_dict, dict) else load_xgboost_dump(path_or_dict)
    learner = doc["learner"]
    objective = (learner.get("objective") or {}).get("name")
    if objective is not None and objective not in _IDENTITY_OBJECTIVES:
        raise ValueError(f"objective {objective!r} transforms the margin; only identity-link regression is supported")
    base = _xgb_scalar(learner["learner_model_param"]["base_score"])
    n_feat = int(_xgb_scalar(learner["learner_model_param"]["num_feature"]))
    if int(_xgb_scalar(learner["learner_model_param"].get("num_target", 1))) != 1:
        raise ValueError("multi-target XGBoost models are not supported")
    booster = learner["gradient_booster"]
    weights = None
    if booster.get("name") == "dart":                    # dart = gbtree + one output weight per tree
        weights = [float(w) for w in booster["weight_drop"]]
        booster = booster["gbtree"]
    elif booster.get("name") not in (None, "gbtree"):
        raise ValueError(f"booster {booster.get('name')!r} is not a tree ensemble")
    trees = []
    for ti, tr in enumerate(booster["model"]["trees"]):
        L, R = tr["left_children"], tr["right_children"]
        cond, feat, dl = tr["split_conditions"], tr["split_indices"], tr["default_left"]
        if any(int(t) != 0 for t in tr.get("split_type", ())):
            raise ValueError(f"tree {ti}: categorical splits are not supported")
        if len(L) > 65535:
            raise ValueError(f"tree {ti}: {len(L)} nodes exceed the format's 65535 per tree")
        w = 1.0 if weights is None else weights[ti]

        def emit(k, new_left, L=L, cond=cond, feat=feat, dl=dl, w=w):
            if L[k] < 0:
                # leaves keep their value in split_conditions; a dart weight multiplies in float32 as XGBoost does
                leaf = np.float32(cond[k]) if w == 1.0 else np.float32(cond[k]) * np.float32(w)
                return (float(leaf), 0, 0, 1)
            if not 0 <= int(feat[k]) < n_feat:
                raise ValueError(f"tree {ti}: split on feature {feat[k]} outside 0..{n_feat - 1}")
            return (float(cond[k]), new_left, int(feat[k]), 2 if dl[k] else 0)
        trees.append(_relabel(L, R, emit))
    return pack_model(trees, base, n_feat)
