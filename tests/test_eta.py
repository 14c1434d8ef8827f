"""ETA: feature rows pinned to the reference's ml.py (golden), the model exporters, and the batched scorer."""
import datetime as dt

import numpy as np
import pytest

from conftest import load_golden
from oracle import oracle as O
from routest_b200 import ml as M


def _encode_case(inp):
    pt = inp["pickup_time"]
    if pt.startswith("datetime("):
        p = dt.datetime(2025, 8, 12, 9, 0, 0)
    else:
        p = dt.datetime.fromisoformat(pt)
    return dict(distance_m=[float(inp["distance_m"] or 0)], weather=[M.encode_weather(inp["weather"])],
                traffic=[M.encode_traffic(inp["traffic"])], weekday=[p.weekday()], hour=[p.hour],
                age=[float(inp.get("driver_age", 30.0) or 30.0)])


GOLD = [c for c in load_golden("eta_feature_cases.json") if isinstance(c["inputs"], dict)]


@pytest.mark.parametrize("case", GOLD, ids=lambda c: str(c["inputs"]["pickup_time"]))
def test_oracle_feature_row_matches_reference(case):
    """or_eta_features == the row the reference hands to model.predict (ml.py:35-51), after the float32 view."""
    assert case["row"]["columns"] == M.FEATURES
    row = O.eta_features(**_encode_case(case["inputs"]))[0]
    assert np.array_equal(row, np.array(case["row"]["values"], dtype=np.float64).astype(np.float32))


def synth_model(seed=0, n_estimators=60, depth=5, n=4000):
    from sklearn.ensemble import GradientBoostingRegressor
    rng = np.random.default_rng(seed)
    cols = dict(distance_m=rng.uniform(200, 40000, n), weather=rng.integers(0, 4, n), traffic=rng.integers(0, 4, n),
                weekday=rng.integers(0, 7, n), hour=rng.integers(0, 24, n), age=rng.uniform(20, 40, n).astype(np.float32))
    X = O.eta_features(**cols)
    y = 4 + X[:, 10] * (2.2 + 1.5 * X[:, 5] + 0.8 * X[:, 4] - 0.3 * X[:, 6]) + 3 * X[:, 1] + \
        2 * np.abs(X[:, 9] - 13) / 6 + rng.normal(0, 1.0, n)
    gbr = GradientBoostingRegressor(n_estimators=n_estimators, max_depth=depth, learning_rate=0.1, random_state=seed)
    gbr.fit(X, y)
    return gbr, M.export_sklearn(gbr)


def synth_pairs(seed, n):
    rng = np.random.default_rng(seed)
    return dict(distance_m=rng.uniform(0, 45000, n), weather=rng.integers(0, 5, n).astype(np.uint8),
                traffic=rng.integers(0, 5, n).astype(np.uint8), weekday=rng.integers(0, 7, n).astype(np.uint8),
                hour=rng.integers(0, 24, n).astype(np.uint8), age=rng.uniform(18, 60, n).astype(np.float32))


def test_sklearn_export_evaluates_like_sklearn():
    gbr, blob = synth_model()
    cols = synth_pairs(1, 5000)
    cols["weather"] %= 4
    cols["traffic"] %= 4
    X = O.eta_features(**cols)
    got = O.eta_eval(blob, X, threads=2)
    want = gbr.predict(X)
    assert np.allclose(got, want, rtol=2e-5, atol=1e-4)  # float32 accumulation vs sklearn's float64


def test_xgboost_json_export():
    """A hand-written dump in XGBoost's JSON schema: x < cond goes left, missing follows default_left."""
    doc = {"learner": {"learner_model_param": {"base_score": "0.5", "num_feature": "12"},
                       "gradient_booster": {"model": {"trees": [
                           {"left_children": [1, -1, 3, -1, -1], "right_children": [2, -1, 4, -1, -1],
                            "split_indices": [10, 0, 5, 0, 0], "split_conditions": [5.0, 1.25, 0.5, -2.0, 7.5],
                            "default_left": [0, 0, 1, 0, 0]},
                           {"left_children": [-1], "right_children": [-1], "split_indices": [0],
                            "split_conditions": [0.125], "default_left": [0]}]}}}}
    blob = M.export_xgboost_json(doc)
    rows = np.zeros((4, 12), np.float32)
    rows[0, 10] = 4.9                      # < 5 -> leaf 1.25
    rows[1, 10] = 5.0                      # not < 5 -> node 2: traffic_Jam 0 < 0.5 -> -2.0
    rows[2, 10], rows[2, 5] = 9.0, 1.0     # -> 7.5
    rows[3, 10] = np.nan                   # missing at root: default right; node 2 sees 0 -> -2.0
    assert np.allclose(O.eta_eval(blob, rows), [0.5 + 1.25 + 0.125, 0.5 - 2.0 + 0.125, 0.5 + 7.5 + 0.125, 0.5 - 2.0 + 0.125])


def _ubj(v, typed=True):
    """Test-side UBJSON writer in the style XGBoost uses: numeric lists become typed `[$t#L n` containers."""
    import struct

    def ilen(n):
        return b"L" + struct.pack(">q", n)
    if v is None:
        return b"Z"
    if isinstance(v, bool):
        return b"T" if v else b"F"
    if isinstance(v, int):
        return b"l" + struct.pack(">i", v) if -2**31 <= v < 2**31 else b"L" + struct.pack(">q", v)
    if isinstance(v, float):
        return b"d" + struct.pack(">f", v)
    if isinstance(v, str):
        b = v.encode()
        return b"S" + ilen(len(b)) + b
    if isinstance(v, list):
        if typed and v and all(isinstance(x, int) and not isinstance(x, bool) for x in v):
            return b"[$l#" + ilen(len(v)) + struct.pack(f">{len(v)}i", *v)
        if typed and v and all(isinstance(x, float) for x in v):
            return b"[$d#" + ilen(len(v)) + struct.pack(f">{len(v)}f", *v)
        return b"[" + b"".join(_ubj(x, typed) for x in v) + b"]"
    if isinstance(v, dict):
        out = b"{"
        for k, x in v.items():
            kb = k.encode()
            out += b"U" + bytes([len(kb)]) + kb + _ubj(x, typed)
        return out + b"}"
    raise TypeError(type(v))


def _xgb_doc(**learner_extra):
    doc = {"learner": {"learner_model_param": {"base_score": "5E-1", "num_feature": "12", "num_target": "1"},
                       "objective": {"name": "reg:squarederror"},
                       "gradient_booster": {"name": "gbtree", "model": {"trees": [
                           {"left_children": [1, -1, 3, -1, -1], "right_children": [2, -1, 4, -1, -1],
                            "split_indices": [10, 0, 5, 0, 0], "split_conditions": [5.0, 1.25, 0.5, -2.0, 7.5],
                            "default_left": [0, 0, 1, 0, 0], "split_type": [0, 0, 0, 0, 0]},
                           {"left_children": [-1], "right_children": [-1], "split_indices": [0],
                            "split_conditions": [0.125], "default_left": [0], "split_type": [0]}]}}},
           "version": [2, 1, 1]}
    doc["learner"].update(learner_extra)
    return doc


@pytest.mark.parametrize("typed", [True, False], ids=["typed-arrays", "plain-arrays"])
def test_xgboost_ubj_dump_reads_like_the_json_dump(tmp_path, typed):
    """XGBoost 2.1.1 (requirements.txt:24) writes `.ubj` by default: same schema, binary container."""
    doc = _xgb_doc()
    assert M.read_ubjson(_ubj(doc, typed)) == doc
    (tmp_path / "m.ubj").write_bytes(_ubj(doc, typed))
    (tmp_path / "m.json").write_text(__import__("json").dumps(doc))
    (tmp_path / "m.bin").write_bytes(_ubj(doc, typed))                # no telling extension: sniffed
    want = M.export_xgboost_json(doc)
    for name in ("m.ubj", "m.json", "m.bin"):
        assert M.export_xgboost_json(str(tmp_path / name)) == want
    with pytest.raises(ValueError):
        M.read_ubjson(_ubj(doc, typed)[:-9])


def test_xgboost_export_refuses_what_the_format_cannot_express():
    with pytest.raises(ValueError, match="objective"):
        M.export_xgboost_json(_xgb_doc(objective={"name": "reg:gamma"}))
    doc = _xgb_doc()
    doc["learner"]["gradient_booster"]["model"]["trees"][0]["split_type"][2] = 1
    with pytest.raises(ValueError, match="categorical"):
        M.export_xgboost_json(doc)
    doc = _xgb_doc()
    doc["learner"]["gradient_booster"]["model"]["trees"][0]["split_indices"][0] = 12
    with pytest.raises(ValueError, match="feature 12"):
        M.export_xgboost_json(doc)
    doc = _xgb_doc()
    doc["learner"]["learner_model_param"]["base_score"] = "[5E-1]"      # 3.x spelling
    assert M.export_xgboost_json(doc) == M.export_xgboost_json(_xgb_doc())


def test_xgboost_dart_weights_scale_the_leaves():
    doc = _xgb_doc()
    gb = doc["learner"]["gradient_booster"]
    doc["learner"]["gradient_booster"] = {"name": "dart", "gbtree": gb, "weight_drop": [0.5, 2.0]}
    rows = np.zeros((2, 12), np.float32)
    rows[0, 10] = 4.9
    rows[1, 10], rows[1, 5] = 9.0, 1.0
    got = O.eta_eval(M.export_xgboost_json(doc), rows)
    assert np.allclose(got, [0.5 + 0.5 * 1.25 + 2.0 * 0.125, 0.5 + 0.5 * 7.5 + 2.0 * 0.125])


def test_predict_without_model_returns_none_pair(monkeypatch, tmp_path):
    """ml.py:25-26: no usable model => (None, None), remembered without retrying."""
    monkeypatch.setenv("ETA_MODEL_PATH", str(tmp_path / "missing.rte"))
    M._state.update(engine=None, loaded=None)
    assert M.predict_eta_minutes(weather="Sunny", traffic="Low", distance_m=1.0, pickup_time=None) == (None, None)
    assert isinstance(M._state["loaded"], str) and M._state["loaded"].startswith("ERROR:")
    M._state.update(engine=None, loaded=None)


# ---------------------------------------------------------------------------------------------------------- GPU

@pytest.mark.gpu
@pytest.mark.parametrize("case", GOLD, ids=lambda c: str(c["inputs"]["pickup_time"]))
def test_engine_feature_row_matches_reference(engine, case):
    row = engine.eta_features(**_encode_case(case["inputs"]))[0]
    assert np.array_equal(row, np.array(case["row"]["values"], dtype=np.float64).astype(np.float32))


@pytest.mark.gpu
def test_engine_scores_equal_oracle_bitwise_1m_pairs(engine):
    """BASELINE config 3: 1M synthetic stop pairs; fused feature build + ensemble, bit-exact vs the scalar oracle
    (north-star bound: 1e-4 relative)."""
    _, blob = synth_model(seed=3, n_estimators=120, depth=6)
    engine.eta_load_model(blob)
    nt, nn = engine.eta_model_info()
    assert nt == 120 and nn > 120
    cols = synth_pairs(3456, 1_000_000)
    cols["age"][::97] = 0.0          # driver_age 0 -> 30 (ml.py:47)
    cols["distance_m"][::89] = np.nan  # NaN stays NaN and follows default directions
    got = engine.eta_score(**cols)
    want = O.eta_eval(blob, O.eta_features(**cols), threads=O.num_threads())
    assert np.array_equal(got, want)
    assert np.array_equal(engine.eta_features(**{k: v[:4096] for k, v in cols.items()}),
                          O.eta_features(**{k: v[:4096] for k, v in cols.items()}), equal_nan=True)


def synth_blob_complete(trees, depth, seed=3456):
    """The bench's ensemble (bench.eta_model_blob): complete trees with random splits, as surveyed for config 3."""
    import bench
    return bench.eta_model_blob(trees, depth, seed)


def test_bench_ensemble_is_a_valid_model_for_the_oracle():
    blob = synth_blob_complete(8, 3)
    cols = synth_pairs(5, 257)
    out = O.eta_eval(blob, O.eta_features(**cols))
    assert out.shape == (257,) and np.isfinite(out).all() and np.unique(out).size > 8


@pytest.mark.gpu
def test_config3_as_surveyed_500_trees_depth6_1m_pairs(engine):
    """1M pairs on a 500-tree depth-6 ensemble (the one bench.py times), bit-exact vs the oracle."""
    blob = synth_blob_complete(500, 6)
    engine.eta_load_model(blob)
    assert engine.eta_model_info() == (500, 500 * 127)
    cols = synth_pairs(3456, 1_000_000)
    got = engine.eta_score(**cols)
    want = O.eta_eval(blob, O.eta_features(**cols), threads=O.num_threads())
    assert np.array_equal(got, want)


@pytest.mark.gpu
def test_eta_pair_ranges_written_by_shards_equal_one_call(engine):
    """SURVEY.md §8(e) row 4 on one GPU: eight contiguous pair ranges scored one after the other == one call."""
    import torch
    from routest_b200.dist import eta_score_shard_dev
    blob = synth_blob_complete(60, 5)
    engine.eta_load_model(blob)
    n = 100_003
    cols = synth_pairs(17, n)
    want = engine.eta_score(**cols)
    dev = [torch.tensor(cols[k], device="cuda") for k in ("distance_m", "weather", "traffic", "weekday", "hour", "age")]
    out = torch.full((n,), -1.0, dtype=torch.float32, device="cuda")
    cuts = [eta_score_shard_dev(engine, [d.data_ptr() for d in dev], n, out.data_ptr(), r, 8,
                                torch.cuda.current_stream().cuda_stream) for r in range(8)]
    torch.cuda.synchronize()
    assert cuts[0][0] == 0 and cuts[-1][1] == n and all(a[1] == b[0] for a, b in zip(cuts, cuts[1:]))
    assert np.array_equal(out.cpu().numpy(), want)


@pytest.mark.gpu
def test_engine_many_chunks_and_ragged_sizes(engine):
    """An ensemble larger than one shared-memory chunk, and pair counts around the block size."""
    _, blob = synth_model(seed=4, n_estimators=400, depth=6, n=3000)
    engine.eta_load_model(blob)
    for n in (1, 255, 1024, 1025, 4099):
        cols = synth_pairs(n, n)
        assert np.array_equal(engine.eta_score(**cols), O.eta_eval(blob, O.eta_features(**cols)))
    assert engine.eta_score([], [], [], [], [], []).size == 0


@pytest.mark.gpu
def test_predict_eta_minutes_mirror(engine):
    """Same signature and return convention as the reference (ml.py:23-58), backed by the engine."""
    gbr, blob = synth_model(seed=5)
    M.use_engine(engine, blob)
    minutes, ts = M.predict_eta_minutes(weather="Stormy", traffic="Jam", distance_m=6983.0,
                                        pickup_time="2025-08-17T23:59:59", driver_age=41)
    X = O.eta_features([6983.0], [1], [1], [6], [23], [41.0])
    assert minutes == float(O.eta_eval(blob, X)[0])
    assert ts == (dt.datetime.fromisoformat("2025-08-17T23:59:59") + dt.timedelta(minutes=minutes)).isoformat()
    assert abs(minutes - float(gbr.predict(X)[0])) < 1e-3
    # unknown strings set no column; None distance -> 0; age None -> 30
    m2, _ = M.predict_eta_minutes(weather="Fog", traffic="low", distance_m=None, pickup_time=dt.datetime(2025, 1, 6, 8),
                                  driver_age=None)
    assert m2 == float(O.eta_eval(blob, O.eta_features([0.0], [255], [255], [0], [8], [30.0]))[0])
    M._state.update(engine=None, loaded=None)


@pytest.mark.gpu
def test_scoring_without_model_is_an_error(engine):
    from routest_b200 import Engine, EngineError
    with Engine(0) as fresh:
        with pytest.raises(EngineError) as ei:
            fresh.eta_score([1.0], [0], [0], [0], [0], [30.0])
        assert "no ETA model loaded" in str(ei.value)
        with pytest.raises(EngineError):
            fresh.eta_load_model(b"nope")
