"""GPU parity tests (B200) for the callers either side of the hot path (SURVEY 8f NEXT-3 / NEXT-4), through the C ABI:
the drop-in against the committed outputs of the reference's own numba / OpenCV code (tests/golden/legacy_callers.npz)
and against the numpy restatement on ragged shapes, tile seams and full-size frames.

Bars: uint8 / uint16 / bool work (adaptive threshold, CLAHE) bit-exact; Liu-Shen float64 <= 1e-10 relative (max
norm), float32 <= 1e-4 relative."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN_DIR
from oracle import flow_oracle as fo
from oracle import legacy_oracle as lo
from opticalflow_b200 import _lib, engine, optical_flow

pytestmark = pytest.mark.gpu


@pytest.fixture(scope='module')
def golden():
    return np.load(os.path.join(GOLDEN_DIR, 'legacy_callers.npz'))


def unpack(bits, shape):
    return np.unpackbits(bits)[:int(np.prod(shape))].reshape(shape).astype(bool)


def relmax(a, b):
    return np.abs(np.asarray(a) - np.asarray(b)).max() / max(np.abs(np.asarray(b)).max(), 1e-300)


def texture(T, Nx, Ny, seed, lo_=20.0, hi=220.0):
    import scipy.ndimage
    rng = np.random.default_rng(seed)
    base = scipy.ndimage.gaussian_filter(rng.standard_normal((Nx + 2 * T, Ny + 2 * T)), 2.0)
    base = (base - base.min()) / (base.max() - base.min()) * (hi - lo_) + lo_
    return np.stack([base[T + t:T + t + Nx, T:T + Ny] + 0.3 * t for t in range(T)])


# ---------------------------------------------------------------------------------------------- Liu-Shen
@pytest.mark.parametrize('dtype,tol', [('float64', 1e-10), ('float32', 1e-4)])
def test_liu_shen_matches_reference_golden(cuda_device, golden, dtype, tol):
    z = golden
    dx, dt, alpha, its = z['ls_a_args']
    got = optical_flow.liu_shen_optical_flow_jit(z['ls_a_movie'], dx, dt, alpha, 1.0, z['ls_a_gx'], z['ls_a_gy'], z['ls_a_gr'],
                                                 int(its), 1e-9, True, dtype=dtype)
    for g, key in zip(got[:4], ('vx', 'vy', 'speed', 'rem')):
        assert g.dtype == np.float64 and g.shape == z['ls_a_' + key].shape
        assert relmax(g, z['ls_a_' + key]) < tol, key
    assert got[4] == int(z['ls_a_last'])
    dx, dt, alpha, its = z['ls_b_args']
    guess = np.full(z['ls_b_movie'].shape[1:], 0.1)
    got = optical_flow.liu_shen_optical_flow_jit(z['ls_b_movie'], dx, dt, alpha, 1.0, guess, guess, None, int(its), dtype=dtype)
    for g, key in zip(got[:3], ('vx', 'vy', 'speed')):
        assert relmax(g, z['ls_b_' + key]) < tol, key


@pytest.mark.parametrize('shape', [(2, 2), (3, 5), (5, 7), (33, 130), (17, 128), (40, 257), (65, 260), (129, 132)])
@pytest.mark.parametrize('iterations', [1, 4, 9])
def test_liu_shen_matches_oracle_on_ragged_shapes(cuda_device, shape, iterations):
    rng = np.random.default_rng(shape[0] * 131 + shape[1] + iterations)
    movie = texture(3, shape[0], shape[1], shape[0] + shape[1])
    gx, gy = rng.standard_normal(shape) * 0.2, rng.standard_normal(shape) * 0.2
    want = lo.liu_shen_optical_flow(movie, 2.0, 0.5, 500.0, 1.0, gx, gy, None, iterations)
    got = optical_flow.liu_shen_optical_flow_jit(movie, 2.0, 0.5, 500.0, 1.0, gx, gy, None, iterations, batch_pairs=1 + iterations % 2)
    for g, w in zip(got[:3], want[:3]):
        assert relmax(g, w) < 1e-10


@pytest.mark.parametrize('shape', [(64, 64), (70, 512), (33, 516), (200, 1028)])
@pytest.mark.parametrize('dtype,tol', [('float64', 1e-10), ('float32', 2e-4)])
def test_liu_shen_streamed_and_tiled_sweeps_agree_with_oracle(cuda_device, monkeypatch, shape, dtype, tol):
    """Rows that are 16-byte multiples take the streamed sweep; OFB_LIU_SHEN_TILED forces the tile kernel on them."""
    rng = np.random.default_rng(shape[1])
    movie = texture(4, shape[0], shape[1], shape[0])
    gx, gy = rng.standard_normal(shape) * 0.2, rng.standard_normal(shape) * 0.2
    want = lo.liu_shen_optical_flow(movie, 1.0, 1.0, 800.0, 1.0, gx, gy, None, 5)
    streamed = optical_flow.liu_shen_optical_flow_jit(movie, 1.0, 1.0, 800.0, 1.0, gx, gy, None, 5, dtype=dtype)
    monkeypatch.setenv('OFB_LIU_SHEN_TILED', '1')
    tiled = optical_flow.liu_shen_optical_flow_jit(movie, 1.0, 1.0, 800.0, 1.0, gx, gy, None, 5, dtype=dtype)
    for q in range(3):
        assert relmax(streamed[q], want[q]) < tol and relmax(tiled[q], want[q]) < tol


def test_liu_shen_sweep_is_affine_at_full_size(cuda_device):
    """1024^2 (oracle-free): one sweep is v -> M v + c, so sweep(a) + sweep(b) - sweep(0) == sweep(a + b); float32 agrees."""
    rng = np.random.default_rng(3)
    N = 1024
    movie = texture(2, N, N, 12)
    a, b = rng.standard_normal((2, N, N)) * 0.3, rng.standard_normal((2, N, N)) * 0.3

    def sweep(gx, gy, dtype='float64', its=1):
        return optical_flow.liu_shen_optical_flow_jit(movie, 1.0, 1.0, 2e3, 1.0, gx, gy, None, its, dtype=dtype)

    zero = np.zeros((N, N))
    sa, sb, s0, sab = sweep(a[0], a[1]), sweep(b[0], b[1]), sweep(zero, zero), sweep(a[0] + b[0], a[1] + b[1])
    for q in range(2):
        assert relmax(sa[q] + sb[q] - s0[q], sab[q]) < 1e-11
    f64, f32 = sweep(a[0], a[1], its=20), sweep(a[0], a[1], 'float32', its=20)
    for q in range(3):
        assert relmax(f32[q], f64[q]) < 1e-4
    # a 192^2 corner window of the full frame against the oracle run on a larger crop (20 sweeps reach 20 pixels)
    crop = lo.liu_shen_optical_flow(movie[:, :256, :256], 1.0, 1.0, 2e3, 1.0, a[0, :256, :256], a[1, :256, :256], None, 20)
    for q in range(3):
        assert relmax(f64[q][0, :192, :192], crop[q][0, :192, :192]) < 1e-10


def test_deprecated_driver_records_iterates(cuda_device):
    movie = texture(3, 24, 36, 4)
    kw = dict(delta_x=0.5, delta_t=2.0, speed_alpha=300.0, use_liu_shen=True)
    res = optical_flow.conduct_variational_optical_flow_deprecated(movie, max_iterations=6, return_iterations=True,
                                                                   iteration_stepsize=2, **kw)
    assert res['v_x_steps'].shape == (2, 4, 24, 36) and res['total_iterations'] == 6 and res['iteration_stepsize'] == 2
    assert np.allclose(res['v_x_steps'][:, 0], 0.1) and np.allclose(res['remodelling'], 0.5)
    for k, its in ((1, 2), (2, 4), (3, 6)):
        direct = optical_flow.conduct_variational_optical_flow_deprecated(movie, max_iterations=its, **kw)
        for key in ('v_x', 'v_y', 'speed'):
            assert relmax(res[key + '_steps'][:, k], direct[key]) < 1e-10
    want = lo.liu_shen_optical_flow(movie, 0.5, 2.0, 300.0, 1.0, np.full((24, 36), 0.1), np.full((24, 36), 0.1), None, 6)
    assert relmax(res['v_x'], want[0]) < 1e-10 and relmax(res['speed'], want[2]) < 1e-10
    blurred = optical_flow.conduct_variational_optical_flow_deprecated(movie, max_iterations=3, smoothing_sigma=1.5, **kw)
    want = lo.liu_shen_optical_flow(fo.blur_movie(movie, 1.5), 0.5, 2.0, 300.0, 1.0, np.full((24, 36), 0.1),
                                    np.full((24, 36), 0.1), None, 3)
    assert relmax(blurred['v_y'], want[1]) < 1e-9
    with pytest.raises(ValueError):
        optical_flow.conduct_variational_optical_flow_deprecated(movie, use_liu_shen=False)


# ---------------------------------------------------------------------------------------------- adaptive threshold
@pytest.mark.parametrize('case', ['th_a', 'th_b', 'th_c'])
def test_adaptive_threshold_matches_reference_golden(cuda_device, golden, case):
    movie = golden[case + '_movie']
    window, threshold = golden[case + '_args']
    got = optical_flow.apply_adaptive_threshold(movie, int(window), threshold)
    assert got.dtype == bool and np.array_equal(got, unpack(golden[case + '_out'], movie.shape))


@pytest.mark.parametrize('shape,window,threshold', [((1, 5, 5), 3, 0.0), ((2, 30, 200), 51, -5), ((1, 64, 64), 255, 0.5),
                                                    ((3, 129, 65), 7, 3.0), ((1, 1024, 1024), 151, -5), ((1, 300, 257), 101, -0.2)])
def test_adaptive_threshold_matches_oracle(cuda_device, shape, window, threshold):
    rng = np.random.default_rng(shape[1] + window)
    movie = np.abs(texture(shape[0], shape[1], shape[2], window)) + rng.uniform(0, 30, shape)
    assert np.array_equal(optical_flow.apply_adaptive_threshold(movie, window, threshold),
                          lo.apply_adaptive_threshold(movie, window, threshold))
    ints = (movie * 10).astype(np.uint16)                      # integer movies divide in float64 like numpy
    assert np.array_equal(optical_flow.apply_adaptive_threshold(ints, window, threshold),
                          lo.apply_adaptive_threshold(ints, window, threshold))


def test_adaptive_threshold_rejects_bad_windows(cuda_device):
    movie = np.ones((1, 16, 16))
    for window in (4, 1, 257):
        with pytest.raises(_lib.OfbError):
            optical_flow.apply_adaptive_threshold(movie, window)


# ---------------------------------------------------------------------------------------------- CLAHE
@pytest.mark.parametrize('case', ['cl_a', 'cl_b', 'cl_c'])
def test_clahe_matches_reference_golden(cuda_device, golden, case):
    movie = golden[case + '_movie']
    clip, tiles = golden[case + '_args']
    got = optical_flow.apply_clahe(movie, clip, int(tiles))
    assert got.dtype == np.float64 and np.array_equal(got, golden[case + '_out'].astype(np.float64))


@pytest.mark.parametrize('shape,clip,tiles,scale', [((1, 16, 16), 40.0, 2, 1.0), ((3, 130, 75), 500.0, 5, 20.0),
                                                    ((2, 64, 250), 50000, 3, 250.0), ((1, 1024, 1024), 10000, 10, 100.0),
                                                    ((1, 101, 103), 3.0, 7, 0.2), ((2, 48, 64), 0.0, 4, 30.0),
                                                    ((1, 40, 40), 50000, 1, 300.0)])
def test_clahe_matches_oracle(cuda_device, shape, clip, tiles, scale):
    movie = texture(shape[0], shape[1], shape[2], tiles) * scale
    want = lo.apply_clahe(movie, clip, tiles)
    assert np.array_equal(optical_flow.apply_clahe(movie, clip, tiles), want)
    # fewer frames in flight than frames: the chunk loop inside ofb_clahe
    d = torch.from_numpy(movie).to(cuda_device)
    tiles_y = int(round(tiles * shape[2] / shape[1]))
    got = engine.clahe(d, clip, tiles, tiles_y, frames_in_flight=1).cpu().numpy()
    assert np.array_equal(got, want)
    got32 = engine.clahe(d.float(), clip, tiles, tiles_y).cpu().numpy()       # float32 frames truncate to the same uint16
    assert np.array_equal(got32, lo.apply_clahe(movie.astype(np.float32), clip, tiles))


# ---------------------------------------------------------------------------------------------- background
def test_background_subtraction_matches_reference_golden(cuda_device, golden):
    got = optical_flow.subtract_background(golden['bg_movie'], float(golden['bg_level']))
    want = golden['bg_out']
    # the mask comes from a blur that agrees to rounding: pixels whose blurred value sits within 1e-9 of the level may flip
    blurred = fo.blur_movie(golden['bg_movie'], 10)
    decided = np.abs(blurred - float(golden['bg_level'])) > 1e-9
    assert decided.mean() > 0.999 and np.abs(got - want)[decided].max() < 1e-10


def test_movie_max(cuda_device):
    rng = np.random.default_rng(1)
    for dtype in (np.float64, np.float32):
        movie = rng.standard_normal((3, 77, 129)).astype(dtype) - 5.0          # all negative: the -inf start matters
        got = engine.movie_max(torch.from_numpy(movie).to(cuda_device))
        assert float(got.cpu()[0]) == float(movie.max())
