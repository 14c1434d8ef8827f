"""GPU: rt_matrix_cells_static_dev — the matrix build for callers that cannot allocate (their GPU memory belongs to
a pool, or cudaMalloc's device-wide synchronisation must stay off their request path).  Its scratch is a static device
array inside the library, so the test checks what such a caller relies on: the same cells as the ordinary entry
points (rt_unit_vectors_dev + rt_matrix_cells_dev into caller buffers), bit for bit as everywhere in this suite, on
two calls in a row — the second with fewer nodes, so stale vectors of the first call sit in the shared array — the
same cells as the CPU oracle computes from the coordinates, and a clear refusal beyond the static size.  Collected last on purpose: a failure here cannot hide any other test."""
import numpy as np
import pytest

from conftest import TS_CAR, synth_instance

pytestmark = pytest.mark.gpu


def _both_paths(engine, torch, lat, lon):
    n = lat.size
    s = torch.cuda.current_stream().cuda_stream
    lat_d, lon_d = torch.tensor(lat, device="cuda"), torch.tensor(lon, device="cuda")
    pitch = engine.matrix_pitch(n)
    # existing entry points: unit vectors into a buffer the caller allocates, then the cells
    xyz = torch.empty((3, n), dtype=torch.float64, device="cuda")
    want = torch.full((n, pitch, 2), -1, dtype=torch.int32, device="cuda")
    engine.unit_vectors_dev(lat_d.data_ptr(), lon_d.data_ptr(), n, 1, xyz.data_ptr(), s)
    engine.matrix_cells_dev(xyz.data_ptr(), n, 1, want.data_ptr(), dist_scale=10.0, time_scale=TS_CAR, stream=s)
    # new entry point: inputs and output only
    got = torch.full((n, pitch, 2), -2, dtype=torch.int32, device="cuda")
    engine.matrix_cells_static_dev(lat_d.data_ptr(), lon_d.data_ptr(), n, got.data_ptr(), dist_scale=10.0,
                                   time_scale=TS_CAR, stream=s)
    torch.cuda.synchronize()
    return got.cpu().numpy(), want.cpu().numpy()


def test_static_scratch_matrix_equals_the_allocating_path_twice_in_a_row(engine):
    import torch
    for seed, stops in ((4242, 1500), (4243, 333)):  # second call is smaller: stale entries stay behind it
        lat, lon, _, _ = synth_instance(seed, stops, 10)
        got, want = _both_paths(engine, torch, lat, lon)
        assert np.array_equal(got, want)
        assert got[:, :stops + 1, 0].any() and not got[:, stops + 1:, :].any()
        # and directly against the oracle (not only against the engine's other path)
        from oracle import oracle as O
        ref = O.matrix_rows(O.unit_vectors(lat, lon), 0, stops + 1, 10.0, TS_CAR, threads=O.num_threads())
        assert np.array_equal(got[:, :stops + 1, 0], ref["dist"]) and np.array_equal(got[:, :stops + 1, 1], ref["time"])


def test_static_scratch_refuses_sizes_beyond_its_array(engine):
    import torch
    from routest_b200 import EngineError
    n = engine.static_max_nodes() + 1
    assert n == 16385
    buf = torch.zeros(16, dtype=torch.float64, device="cuda")  # never touched: the size check comes first
    with pytest.raises(EngineError) as ei:
        engine.matrix_cells_static_dev(buf.data_ptr(), buf.data_ptr(), n, buf.data_ptr())
    assert "exceeds the static scratch" in str(ei.value)
