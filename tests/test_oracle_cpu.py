"""CPU tests: the oracle against the committed golden vectors (outputs of the unmodified reference) and,
when /root/reference is present, against the reference itself; the numpy solver model; host helpers."""
import glob
import os

import numpy as np
import pytest

from conftest import GOLDEN_DIR
from oracle import flow_oracle as fo
from oracle import mg_model as mg
from oracle import ref_import
from opticalflow_b200 import synthetic


def load_golden(name):
    z = np.load(os.path.join(GOLDEN_DIR, name + '.npz'))
    kwargs = {}
    for key in z.files:
        if key.startswith('kw_'):
            v = float(z[key])
            kwargs[key[3:]] = None if np.isnan(v) else v
    return z, kwargs


GOLDEN_NAMES = sorted(os.path.splitext(os.path.basename(p))[0] for p in glob.glob(os.path.join(GOLDEN_DIR, '*.npz'))
                      if not os.path.basename(p).startswith('legacy_')        # legacy_*.npz: tests/test_legacy_*.py
                      and not os.path.basename(p).endswith('_sample.npz'))    # samples of bench-size solves: -m gpu tests


def test_golden_fixtures_exist():
    assert {'hat50', 'texture64', 'ragged_40x56', 'texture96_realregime', 'tiny_5x7'} <= set(GOLDEN_NAMES)


@pytest.mark.parametrize('name', GOLDEN_NAMES)
def test_oracle_matches_reference_golden(name):
    z, kwargs = load_golden(name)
    got = fo.variational_optical_flow(z['movie'], **kwargs)
    for key in ('v_x', 'v_y', 'speed', 'remodelling'):
        scale = max(np.linalg.norm(z[key]), 1e-300)
        assert np.linalg.norm(got[key] - z[key]) / scale < 1e-9, key
    want_fun = z['functionals']
    got_fun = np.array([got['L1_functional'], got['remodelling_functional'], got['speed_functional']])
    np.testing.assert_allclose(got_fun, want_fun, rtol=1e-8)
    if 'blurred_data' in z.files:
        np.testing.assert_allclose(got['blurred_data'], z['blurred_data'], rtol=0, atol=1e-10)


def test_hat_demo_known_means():
    """Means the survey probed from the reference for simple_test_with_data_on_boundary."""
    z, kwargs = load_golden('hat50')
    assert abs(z['v_x'].mean() - 0.1049781) < 1e-6
    assert abs(z['v_y'].mean() - 0.1950244) < 1e-6
    assert abs(z['remodelling'].mean() - 0.0486220) < 1e-6


@pytest.mark.skipif(not ref_import.reference_available(), reason='reference sources only exist in the build container')
def test_oracle_operator_equals_reference_matrix():
    rng = np.random.default_rng(5)
    P = rng.uniform(20, 200, (23, 31))
    C = P + rng.normal(0, 3, P.shape)
    A, b = ref_import.reference_system(P, C, 123.0, 45.0)
    assert abs(A - fo.assemble_csr(P, 123.0, 45.0)).max() == 0.0
    x = rng.standard_normal((3, 23, 31))
    y = fo.planes_to_interleaved(fo.apply_operator(P, x, 123.0, 45.0))
    yr = A @ fo.planes_to_interleaved(x)
    assert np.linalg.norm(y - yr) / np.linalg.norm(yr) < 1e-14
    assert np.abs(fo.planes_to_interleaved(fo.build_rhs(P, C)) - b).max() == 0.0


def test_blur_matches_scipy_ndimage():
    import scipy.ndimage
    rng = np.random.default_rng(1)
    f = rng.uniform(0, 255, (37, 61))
    for sigma in (0.7, 2.5, 5.0):
        want = scipy.ndimage.gaussian_filter(f, sigma, mode='nearest', truncate=4.0)
        assert np.abs(fo.gaussian_blur(f, sigma) - want).max() < 1e-11


def test_eliminated_operator_and_mg_model_solve():
    """The eliminated system + ghost completion reproduces the direct solution of the full system."""
    movie = fo.blur_movie(synthetic.translating_movie(32, 40, 2, seed=4), 2.0)
    P, C = movie
    alpha, beta = 1e5, 1e4
    b = mg.zero_ghosts(fo.build_rhs(P, C))
    H = mg.Hierarchy(P, alpha, beta)
    x, it, hist = mg.bicgstab(lambda v: H.apply(0, v), b, np.zeros_like(b), lambda r: H.jacobi(0, r),
                              rtol=1e-11, max_it=4000)
    xd, _ = fo.solve_pair(P, C, alpha, beta)
    xe = mg.fill_ghosts(x.copy())
    assert np.linalg.norm(xe - xd) / np.linalg.norm(xd) < 1e-7


def test_host_helpers_match_oracle():
    from opticalflow_b200 import optical_flow as of
    rng = np.random.default_rng(3)
    m = rng.standard_normal((9, 12))
    for rule in ('dx', 'dy', 'dxy', 'dyx', 'dxx', 'dyy', 'bar_x', 'bar_y', 'bar'):
        np.testing.assert_allclose(of.apply_numerical_derivative(m, rule), fo.numerical_derivative(m, rule), atol=1e-15)
    a, b = m.copy(), m.copy()
    of.apply_constant_boundary_condition(a)
    fo.apply_constant_boundary_condition(b)
    assert np.array_equal(a, b)
    for q in ('ux', 'uy', 'remodelling'):
        for inc in (False, True):
            assert np.array_equal(of.get_index_set(7, 9, 0, 0, q, inc), fo.get_index_set(7, 9, 0, 0, q, inc))
        assert np.array_equal(of.get_index_set(7, 9, -1, 1, q), fo.get_index_set(7, 9, -1, 1, q))
    with pytest.raises(ValueError):
        of.get_index_set(7, 9, 0, 0, 'bogus')
    assert of.format_elapsed_time(125.25) == (2, 5, 250)
