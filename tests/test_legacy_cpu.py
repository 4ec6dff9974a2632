"""CPU tests for the callers either side of the hot path (SURVEY 8f NEXT-3 / NEXT-4): the numpy restatement
(oracle/legacy_oracle.py) against the committed outputs of the reference's own numba / OpenCV code
(tests/golden/legacy_callers.npz, oracle/make_golden.py), and against the reference itself where it is present."""
import os

import numpy as np
import pytest

from conftest import GOLDEN_DIR
from oracle import legacy_oracle as lo
from oracle import ref_import


@pytest.fixture(scope='module')
def golden():
    return np.load(os.path.join(GOLDEN_DIR, 'legacy_callers.npz'))


def unpack(bits, shape):
    return np.unpackbits(bits)[:int(np.prod(shape))].reshape(shape).astype(bool)


def test_liu_shen_oracle_matches_reference_golden(golden):
    z = golden
    dx, dt, alpha, its = z['ls_a_args']
    got = lo.liu_shen_optical_flow(z['ls_a_movie'], dx, dt, alpha, 1.0, z['ls_a_gx'], z['ls_a_gy'], z['ls_a_gr'], int(its))
    for g, key in zip(got[:4], ('vx', 'vy', 'speed', 'rem')):
        assert np.abs(g - z['ls_a_' + key]).max() <= 1e-12 * max(1.0, np.abs(z['ls_a_' + key]).max()), key
    assert got[4] == int(z['ls_a_last'])
    dx, dt, alpha, its = z['ls_b_args']
    guess = np.full(z['ls_b_movie'].shape[1:], 0.1)
    got = lo.liu_shen_optical_flow(z['ls_b_movie'], dx, dt, alpha, 1.0, guess, guess, None, int(its))
    for g, key in zip(got[:3], ('vx', 'vy', 'speed')):
        assert np.abs(g - z['ls_b_' + key]).max() <= 1e-12 * max(1.0, np.abs(z['ls_b_' + key]).max()), key


@pytest.mark.parametrize('case', ['th_a', 'th_b', 'th_c'])
def test_adaptive_threshold_oracle_is_bit_exact(golden, case):
    movie = golden[case + '_movie']
    window, threshold = golden[case + '_args']
    got = lo.apply_adaptive_threshold(movie, int(window), threshold)
    want = unpack(golden[case + '_out'], movie.shape)
    assert got.dtype == bool and np.array_equal(got, want)
    assert 0.05 < want.mean() < 0.95            # the fixture exercises both outcomes


@pytest.mark.parametrize('case', ['cl_a', 'cl_b', 'cl_c'])
def test_clahe_oracle_is_bit_exact(golden, case):
    movie = golden[case + '_movie']
    clip, tiles = golden[case + '_args']
    got = lo.apply_clahe(movie, clip, int(tiles))
    assert got.dtype == np.float64 and np.array_equal(got, golden[case + '_out'].astype(np.float64))


def test_background_subtraction_oracle(golden):
    from oracle import flow_oracle as fo
    movie, level = golden['bg_movie'], float(golden['bg_level'])
    got = lo.subtract_background(movie, fo.blur_movie(movie, 10), level)
    assert np.abs(got - golden['bg_out']).max() < 1e-10
    assert 0.05 < (golden['bg_out'] == 0).mean() < 0.95


def test_clahe_geometry_follows_opencv_size_convention():
    # tileGridSize=(tile_number, round(tile_number*Ny/Nx)) is cv::Size(width, height): the first splits the columns
    assert lo.clahe_geometry(100, 200, 10)[:2] == (10, 20)
    assert lo.clahe_geometry(90, 70, 10) == (10, 8, 96, 80, 8, 12)


@pytest.mark.skipif(not ref_import.reference_available(), reason='reference sources not present (GPU box)')
def test_oracle_against_reference_run_here():
    """Fresh inputs (not the committed ones) through the unmodified reference's OpenCV paths."""
    import contextlib
    import io
    ref = ref_import.import_reference()
    rng = np.random.default_rng(5)
    movie = rng.uniform(0, 4000, (2, 75, 85))
    assert np.array_equal(ref.apply_adaptive_threshold(movie, 21, 1.5), lo.apply_adaptive_threshold(movie, 21, 1.5))
    with contextlib.redirect_stdout(io.StringIO()):
        want = ref.apply_clahe(movie, 3000, 6)
    assert np.array_equal(want, lo.apply_clahe(movie, 3000, 6))
    # edge cases: a window larger than the frame, no clipping (clipLimit 0), one tile, few grey levels
    small = rng.uniform(0, 255, (1, 20, 23))
    assert np.array_equal(ref.apply_adaptive_threshold(small, 51, -2.0), lo.apply_adaptive_threshold(small, 51, -2.0))
    flat = np.round(rng.uniform(0, 6, (1, 48, 64))) * 100.0
    for clip, tiles in ((0.0, 4), (50000, 1), (1.0, 3)):
        with contextlib.redirect_stdout(io.StringIO()):
            want = ref.apply_clahe(flat, clip, tiles)
        assert np.array_equal(want, lo.apply_clahe(flat, clip, tiles)), (clip, tiles)


@pytest.mark.skipif(not ref_import.reference_available(), reason='reference sources not present (GPU box)')
def test_liu_shen_oracle_against_reference_numba_on_tiny_frames():
    """The smallest frames the padded reference arrays allow (2x2, 2x5, 3x3): every pixel is an edge or corner pixel."""
    ref = ref_import.import_reference()
    rng = np.random.default_rng(9)
    for shape, its in (((2, 2), 1), ((2, 5), 3), ((3, 3), 4)):
        movie = rng.uniform(20, 200, (3,) + shape)
        gx, gy, gr = rng.standard_normal(shape), rng.standard_normal(shape), rng.standard_normal(shape)
        want = ref.liu_shen_optical_flow_jit(movie, 1.5, 0.5, 50.0, 1.0, gx, gy, gr, its, 1e-9, True)
        got = lo.liu_shen_optical_flow(movie, 1.5, 0.5, 50.0, 1.0, gx, gy, gr, its)
        for g, w in zip(got[:4], want[:4]):
            assert np.abs(g - w).max() <= 1e-12 * max(1.0, np.abs(w).max()), shape
        assert got[4] == want[4]
