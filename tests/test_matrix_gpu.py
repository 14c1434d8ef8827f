"""GPU: matrix build through the C ABI against the oracle (SPEC §1-§2), bit-exact for the integer planes."""
import numpy as np
import pytest

from conftest import synth_instance
from oracle import oracle as O

pytestmark = pytest.mark.gpu

TS = 3.6 / 40.0  # driving-car seconds per metre


def _oracle(lat, lon, r0, r1, ds, ts, want):
    return O.matrix_rows(O.unit_vectors(lat, lon), r0, r1, ds, ts, want=want, threads=O.num_threads())


@pytest.mark.parametrize("seed,stops", [(1234, 100), (2345, 1000), (11, 1), (12, 15), (13, 16), (14, 31), (15, 1023), (16, 1024), (17, 2049)])
def test_matrix_bit_exact_metro(engine, seed, stops):
    lat, lon, _, _ = synth_instance(seed, stops, 10)
    n = stops + 1
    got = engine.matrix_build(lat, lon, dist_scale=10.0, time_scale=TS, want=("dist", "time", "metres"))
    ref = _oracle(lat, lon, 0, n, 10.0, TS, ("dist", "time", "metres"))
    assert np.array_equal(got["dist"], ref["dist"])
    assert np.array_equal(got["time"], ref["time"])
    assert np.array_equal(got["metres"], ref["metres"])  # same FP64 operation sequence => identical bits


def test_matrix_vs_libm_haversine(engine):
    """North-star bound: matrix values within 1e-6 relative of the great-circle distance."""
    lat, lon, _, _ = synth_instance(2345, 1000, 10)
    got = engine.matrix_build(lat, lon, dist_scale=10.0, time_scale=TS, want=("dist", "metres"))
    h = O.haversine_libm(lat, lon)
    off = h > 0
    assert (np.abs(got["metres"] - h)[off] / h[off]).max() < 1e-6
    assert (np.abs(got["dist"] - 10.0 * h) <= 0.5 + 1e-6 * 10.0 * h).all()


def test_matrix_global_far_pairs(engine):
    """Half-angle path (k > 0): antipodes, poles, the date line."""
    rng = np.random.default_rng(7)
    lat = np.concatenate([rng.uniform(-90, 90, 300), [90.0, -90.0, 0.0, 0.0, 0.0, 14.5, -14.5]])
    lon = np.concatenate([rng.uniform(-180, 180, 300), [0.0, 0.0, 180.0, -180.0, 0.0, 121.0, -59.0]])
    got = engine.matrix_build(lat, lon, dist_scale=0.01, time_scale=0.001, want=("dist", "time", "metres"))
    ref = _oracle(lat, lon, 0, lat.size, 0.01, 0.001, ("dist", "time", "metres"))
    for k in ("dist", "time", "metres"):
        assert np.array_equal(got[k], ref[k]), k


def test_matrix_row_shard(engine):
    """Row-sharded build (what each rank does at N>1): rows [r0,r1) equal the same rows of the full matrix."""
    lat, lon, _, _ = synth_instance(4567, 700, 10)
    n = 701
    full = _oracle(lat, lon, 0, n, 10.0, TS, ("dist", "time"))
    for r0, r1 in ((0, 88), (88, 350), (350, 701), (700, 701), (5, 5)):
        got = engine.matrix_build(lat, lon, dist_scale=10.0, time_scale=TS, rows=(r0, r1))
        assert np.array_equal(got["dist"], full["dist"][r0:r1])
        assert np.array_equal(got["time"], full["time"][r0:r1])


def test_matrix_single_plane_outputs(engine):
    lat, lon, _, _ = synth_instance(5, 200, 10)
    ref = _oracle(lat, lon, 0, 201, 10.0, TS, ("dist", "time"))
    assert np.array_equal(engine.matrix_build(lat, lon, time_scale=TS, want=("dist",))["dist"], ref["dist"])
    assert np.array_equal(engine.matrix_build(lat, lon, time_scale=TS, want=("time",))["time"], ref["time"])


def test_matrix_chunked_rows_large(engine):
    """5,001 nodes forces more than one row chunk through the double-buffered D2H path; compare by row samples
    plus exact symmetry and a zero diagonal (size-independent properties)."""
    lat, lon, _, _ = synth_instance(5678, 9000, 50)
    n = 9001
    got = engine.matrix_build(lat, lon, dist_scale=10.0, time_scale=TS)
    d, t = got["dist"], got["time"]
    assert np.array_equal(d, d.T) and np.array_equal(t, t.T)
    assert not d.diagonal().any() and not t.diagonal().any()
    xyz = O.unit_vectors(lat, lon)
    for r0 in (0, 3354, 3355, 6710, 8990):
        ref = O.matrix_rows(xyz, r0, min(n, r0 + 11), 10.0, TS)
        assert np.array_equal(d[r0:r0 + 11], ref["dist"]) and np.array_equal(t[r0:r0 + 11], ref["time"])


def test_matrix_cells_dev_batched_layout(engine):
    """Device-resident interleaved layout ({dist,time} int2, pitch 16 cells, zero pad) for a batch of instances."""
    import torch
    B, stops = 3, 77
    n = stops + 1
    lats, lons = [], []
    for b in range(B):
        la, lo, _, _ = synth_instance(100 + b, stops, 5)
        lats.append(la)
        lons.append(lo)
    lat_d = torch.tensor(np.stack(lats), device="cuda")
    lon_d = torch.tensor(np.stack(lons), device="cuda")
    xyz_d = torch.empty((B, 3, n), dtype=torch.float64, device="cuda")
    pitch = engine.matrix_pitch(n)
    cells = torch.full((B, n, pitch, 2), -1, dtype=torch.int32, device="cuda")
    s = torch.cuda.current_stream().cuda_stream
    engine.unit_vectors_dev(lat_d.data_ptr(), lon_d.data_ptr(), n, B, xyz_d.data_ptr(), s)
    engine.matrix_cells_dev(xyz_d.data_ptr(), n, B, cells.data_ptr(), dist_scale=10.0, time_scale=TS, stream=s)
    torch.cuda.synchronize()
    got = cells.cpu().numpy()
    xyz_h = xyz_d.cpu().numpy()
    for b in range(B):
        assert np.array_equal(xyz_h[b], O.unit_vectors(lats[b], lons[b]))
        ref = O.matrix_rows(xyz_h[b], 0, n, 10.0, TS)
        assert np.array_equal(got[b, :, :n, 0], ref["dist"]) and np.array_equal(got[b, :, :n, 1], ref["time"])
        assert not got[b, :, n:, :].any()


def test_matrix_rejects_bad_input(engine):
    from routest_b200 import EngineError
    with pytest.raises(EngineError):
        engine.matrix_build([91.0, 0.0], [0.0, 0.0])
    with pytest.raises(EngineError):
        engine.matrix_build([0.0, float("nan")], [0.0, 0.0])
    with pytest.raises(EngineError):
        engine.matrix_build([0.0, 1.0], [0.0, 1.0], dist_scale=1e6)


def test_fast_sqrt_is_ieee_on_the_whole_admitted_range(engine):
    """The straight-line cell uses nvcc's own post-range-check sqrt sequence; it must equal sqrtf bit for bit for
    EVERY float the fast path admits: [2^-100, 0.0625] (checked up to 4.0 for margin)."""
    lo = int(np.float32(2.0 ** -100).view(np.uint32))
    hi = int(np.float32(4.0).view(np.uint32)) + 1
    assert engine.selftest_fast_sqrt(lo, hi) == 0


def test_closing_the_engine_closes_its_problems_first():
    """A problem's handle points into its engine: Engine.close() destroys the problems it handed out, and a call on
    a closed problem is an error, not a crash."""
    from routest_b200 import Engine, EngineError
    eng = Engine(0)
    pr = eng.problem(11, 2)
    rng = np.random.default_rng(0)
    pr.set_coords(rng.uniform(14.4, 14.6, (2, 11)), rng.uniform(120.9, 121.1, (2, 11)))
    eng.close()
    assert pr.handle is None
    with pytest.raises(EngineError):
        pr.build_matrix(10.0, 0.09)
    pr.close()
    eng.close()
