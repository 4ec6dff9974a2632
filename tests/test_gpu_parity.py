"""GPU parity tests (B200): every stage through the C ABI against the oracle, the drop-in against the committed
golden vectors of the unmodified reference, and size-independent properties at 1024^2.

Tolerances (BASELINE.json north_star): fp64 <= 1e-6 relative L2 on returned fields; fp32 <= 1e-3 relative L2
and <= 1e-2 px max abs on v_x, v_y."""
import numpy as np
import pytest
import torch

from oracle import flow_oracle as fo
from oracle import mg_model as mg
from opticalflow_b200 import engine, optical_flow, synthetic
from test_oracle_cpu import GOLDEN_NAMES, load_golden

pytestmark = pytest.mark.gpu


def rel(a, b):
    return np.linalg.norm(np.asarray(a) - np.asarray(b)) / max(np.linalg.norm(np.asarray(b)), 1e-300)


def dev_t(a, device, dtype=torch.float64):
    return torch.from_numpy(np.ascontiguousarray(a)).to(device).to(dtype).contiguous()


SHAPES = [(5, 5), (7, 6), (23, 31), (40, 56), (64, 64), (33, 130), (129, 257)]


@pytest.mark.parametrize('shape', SHAPES)
@pytest.mark.parametrize('dtype,tol', [(torch.float64, 1e-13), (torch.float32, 2e-5)])
def test_full_operator_matches_reference_matrix(cuda_device, shape, dtype, tol):
    rng = np.random.default_rng(shape[0] * 1000 + shape[1])
    P = rng.uniform(20, 200, shape)
    x = rng.standard_normal((3,) + shape)
    alpha, beta = 123.0, 45.0
    want = fo.apply_operator(P, x, alpha, beta)          # == reference CSR @ x (tests/test_oracle_cpu.py)
    got = engine.apply_operator(dev_t(P[None], cuda_device, dtype), dev_t(x[None], cuda_device, dtype), alpha, beta)
    assert rel(got[0].cpu().double().numpy(), want) < tol


@pytest.mark.parametrize('shape', SHAPES)
@pytest.mark.parametrize('dtype,tol', [(torch.float64, 1e-13), (torch.float32, 2e-5)])
def test_eliminated_operator_matches_model(cuda_device, shape, dtype, tol):
    rng = np.random.default_rng(shape[0] * 77 + shape[1])
    B = 3
    P = rng.uniform(20, 200, (B,) + shape)
    x = rng.standard_normal((B, 3) + shape)
    alphas, betas = np.array([1.0, 1e3, 1e5]), np.array([1e4, 10.0, 1e2])
    got = engine.apply_operator(dev_t(P, cuda_device, dtype), dev_t(x, cuda_device, dtype),
                                torch.from_numpy(alphas), torch.from_numpy(betas), eliminated=True).cpu().double().numpy()
    for k in range(B):
        want = mg.apply_eliminated(P[k], x[k], alphas[k], betas[k])
        assert rel(got[k], want) < tol


# The float32 blur bar: a separable Gaussian of n = 2 int(4 sigma + 0.5) + 1 positive taps summed in float32 carries a relative error
# of at most ~(n + 1) u per axis (u = 6e-8); on data in [0, 255] and both axes that is 255 * 2 * (n + 1) * u: 7e-4 at sigma 2.5 (n = 21),
# 1.7e-3 at sigma 9 (n = 73).  2e-3 absolute is that worst-case bound, 1e-5 of the data range; the float32 flow tolerance (1e-3 relative)
# is two orders of magnitude looser.  Float64 is held to 1e-10.
BLUR32_TOL = 2e-3


@pytest.mark.parametrize('shape', [(5, 5), (23, 31), (64, 64), (129, 257)])
def test_rhs_and_blur(cuda_device, shape):
    rng = np.random.default_rng(9)
    movie = rng.uniform(0, 255, (3,) + shape)
    d = dev_t(movie, cuda_device)
    b = engine.build_rhs(d[:-1].contiguous(), d[1:].contiguous()).cpu().numpy()
    for k in range(2):
        assert rel(b[k], fo.build_rhs(movie[k], movie[k + 1])) < 1e-14
    for sigma in (0.8, 2.5):
        got = engine.blur(d, sigma).cpu().numpy()
        assert np.abs(got - fo.blur_movie(movie, sigma)).max() < 1e-10
        got32 = engine.blur(d.float(), sigma).cpu().numpy()
        assert np.abs(got32 - fo.blur_movie(movie, sigma)).max() < BLUR32_TOL


@pytest.mark.parametrize('shape,sigma', [((150, 300), 5.0), ((70, 131), 9.0), ((96, 260), 17.0), ((1024, 1024), 2.5)])
def test_blur_tiles_and_wide_kernels(cuda_device, shape, sigma):
    """Tile seams, ragged right/bottom tiles, radius > tile height, and the radius > 64 two-launch path."""
    rng = np.random.default_rng(19)
    movie = rng.uniform(0, 255, (2,) + shape)
    want = fo.blur_movie(movie, sigma)
    got = engine.blur(dev_t(movie, cuda_device), sigma).cpu().numpy()
    assert np.abs(got - want).max() < 1e-10
    got32 = engine.blur(dev_t(movie, cuda_device).float(), sigma).cpu().numpy()
    assert np.abs(got32 - want).max() < BLUR32_TOL


def test_multigrid_hierarchy_matches_model(cuda_device):
    """One V(2,2) application on the device == the numpy model of the same cycle (averaged coefficients)."""
    import importlib.util, os, sys
    movie = fo.blur_movie(synthetic.translating_movie(66, 50, 2, seed=5), 2.0)
    P = movie[0]
    rng = np.random.default_rng(2)
    r = mg.zero_ghosts(rng.standard_normal((3,) + P.shape))
    z = engine.precondition(dev_t(P[None], cuda_device), dev_t(r[None], cuda_device), 1e5, 1e4)[0].cpu().numpy()
    # a preconditioner must reduce the residual of S z = r substantially in one application
    res = r - mg.apply_eliminated(P, z, 1e5, 1e4)
    assert np.linalg.norm(res) / np.linalg.norm(r) < 0.35
    assert np.abs(z[:, 0, :]).max() == 0.0 and np.abs(z[:, :, -1]).max() == 0.0


@pytest.mark.parametrize('precond', ['mg', 'bjacobi'])
@pytest.mark.parametrize('shape', [(24, 24), (40, 56), (64, 64)])
def test_solve_fp64_matches_direct_solution(cuda_device, shape, precond):
    movie = fo.blur_movie(synthetic.translating_movie(shape[0], shape[1], 2, seed=6), 2.5)
    P, C = movie
    alpha, beta = 1e5, 1e4
    xd, _ = fo.solve_pair(P, C, alpha, beta)
    Pd, Cd = dev_t(P[None], cuda_device), dev_t(C[None], cuda_device)
    b = engine.build_rhs(Pd, Cd)
    x = torch.zeros_like(b)
    it, rr, cv = engine.solve(Pd, b, x, alpha, beta, rtol=1e-11, max_it=5000, options=engine.make_options(precond))
    assert int(cv[0]) == 1, (int(it[0]), float(rr[0]))
    assert rel(x[0].cpu().numpy(), xd) < 1e-7                       # raw solution, ghost ring included
    if precond == 'mg':
        assert int(it[0]) <= 40


@pytest.mark.parametrize('shape', [(5, 5), (7, 6), (24, 24), (40, 56), (64, 64), (33, 130)])
@pytest.mark.parametrize('alpha', [1e5, 1e3])
def test_direct_solve_matches_superlu(cuda_device, shape, alpha):
    """ofb_direct_solve (multifrontal LU on a nested dissection, fp64) against the oracle's SuperLU solve of the
    reference's matrix (:1146-1147) - also at alpha = 1e3, where the operator is indefinite and the Krylov path fails."""
    movie = fo.blur_movie(synthetic.rotating_shearing_movie(shape[0], shape[1], 3, seed=2), 2.5)
    want = [fo.solve_pair(movie[k], movie[k + 1], alpha, 1e4)[0] for k in range(2)]
    Pd, Cd = dev_t(movie[:-1], cuda_device), dev_t(movie[1:], cuda_device)
    b = engine.build_rhs(Pd, Cd)
    x, rr, cv = engine.direct_solve(Pd, b, alpha, 1e4)
    assert cv.tolist() == [1, 1]
    for k in range(2):
        assert rel(x[k].cpu().numpy(), want[k]) < 1e-9, (k, float(rr[k]))         # raw solution, ghost ring included
        assert float(rr[k]) < 1e-10
    # float32 images and right-hand side, fp64 factorisation: within the float32 tolerance
    x32, rr32, cv32 = engine.direct_solve(Pd.float().contiguous(), b.float().contiguous(), alpha, 1e4)
    assert cv32.tolist() == [1, 1]
    for k in range(2):
        assert rel(x32[k].cpu().double().numpy(), want[k]) < 1e-3


def test_direct_solve_graph_replay_matches_eager(cuda_device, monkeypatch):
    """OFB_DIRECT_GRAPH=1: the launch sequence of a frame shape captured once over staging slots and replayed per pair - the same
    answers as the launch-by-launch path, also when the batch size and the weights change between calls on one workspace."""
    from opticalflow_b200 import _lib
    movie = fo.blur_movie(synthetic.rotating_shearing_movie(40, 56, 4, seed=2), 2.5)
    Pd, Cd = dev_t(movie[:-1], cuda_device), dev_t(movie[1:], cuda_device)
    b = engine.build_rhs(Pd, Cd)
    ws = engine.Workspace()
    monkeypatch.delenv('OFB_DIRECT_GRAPH', raising=False)
    eager = [engine.direct_solve(Pd, b, a, 1e4, ws)[0].clone() for a in (1e5, 1e3)]
    monkeypatch.setenv('OFB_DIRECT_GRAPH', '1')
    for k, a in enumerate((1e5, 1e3)):
        x, rr, cv = engine.direct_solve(Pd, b, a, 1e4, ws)                        # three pairs
        assert _lib.load().ofb_direct_graph_state() == 1 and cv.tolist() == [1, 1, 1]
        assert float((x - eager[k]).abs().max()) <= 1e-9 * float(eager[k].abs().max())
        x1, _, _ = engine.direct_solve(Pd[1:2].contiguous(), b[1:2].contiguous(), a, 1e4, ws)      # one pair, same workspace, same graph
        assert float((x1[0] - eager[k][1]).abs().max()) <= 1e-9 * float(eager[k].abs().max())


def test_solve_initial_guess_and_batch_weights(cuda_device):
    """Per-pair weights in one batch; nonzero initial guess; converged pairs stop iterating."""
    movie = fo.blur_movie(synthetic.translating_movie(48, 48, 3, seed=8), 2.5)
    alphas, betas = np.array([1e5, 2e4]), np.array([1e4, 1e3])
    Pd, Cd = dev_t(movie[:-1], cuda_device), dev_t(movie[1:], cuda_device)
    b = engine.build_rhs(Pd, Cd)
    x = torch.full_like(b, 0.3)
    it, rr, cv = engine.solve(Pd, b, x, torch.from_numpy(alphas), torch.from_numpy(betas), rtol=1e-11, max_it=500)
    assert cv.tolist() == [1, 1]
    for k in range(2):
        xd, _ = fo.solve_pair(movie[k], movie[k + 1], alphas[k], betas[k])
        assert rel(x[k].cpu().numpy(), xd) < 1e-7


@pytest.mark.parametrize('name', GOLDEN_NAMES)
def test_dropin_fp64_matches_reference_golden(cuda_device, name):
    z, kwargs = load_golden(name)
    got = optical_flow.variational_optical_flow(z['movie'], dtype='float64', rtol=1e-11, max_it=20000, **kwargs)
    assert got['converged']
    for key in ('v_x', 'v_y', 'speed', 'remodelling'):
        assert rel(got[key], z[key]) < 1e-6, (key, got['solver_info'])
    fun = np.array([got['L1_functional'], got['remodelling_functional'], got['speed_functional']])
    np.testing.assert_allclose(fun, z['functionals'], rtol=1e-6)
    if 'blurred_data' in z.files:
        assert np.abs(got['blurred_data'] - z['blurred_data']).max() < 1e-9
    assert set(got) - {'solver_info'} == {'v_x', 'v_y', 'speed', 'remodelling', 'original_data', 'delta_x', 'delta_t',
                                         'blurred_data', 'converged', 'L1_functional', 'remodelling_functional',
                                         'speed_functional'}
    assert got['v_x'].dtype == np.float64 and got['v_x'].shape == z['v_x'].shape


@pytest.mark.parametrize('name', ['texture64', 'texture96_realregime', 'ragged_40x56'])
def test_dropin_fp32_within_stated_tolerance(cuda_device, name):
    z, kwargs = load_golden(name)
    got = optical_flow.variational_optical_flow(z['movie'], dtype='float32', rtol=1e-6, **kwargs)
    scale = kwargs.get('delta_x', 1.0) / kwargs.get('delta_t', 1.0)
    for key in ('v_x', 'v_y'):
        assert rel(got[key], z[key]) < 1e-3, (key, got['solver_info'])
        assert np.abs(got[key] - z[key]).max() < 1e-2 * scale
    assert rel(got['remodelling'], z['remodelling']) < 1e-3


def test_dropin_256_config_d1(cuda_device):
    """BASELINE.json configs[0]: 256^2 pair translated by (1.5,-0.7); oracle direct solve takes ~30 s."""
    movie = synthetic.translating_movie(256, 256, 2)
    kw = dict(speed_alpha=1e5, remodelling_alpha=1e4, smoothing_sigma=2.5)
    want = fo.variational_optical_flow(movie, **kw)
    got64 = optical_flow.variational_optical_flow(movie, dtype='float64', rtol=1e-10, **kw)
    got32 = optical_flow.variational_optical_flow(movie, dtype='float32', rtol=1e-6, **kw)
    for key in ('v_x', 'v_y', 'remodelling'):
        assert rel(got64[key], want[key]) < 1e-6, key
        assert rel(got32[key], want[key]) < 1e-3, key
    assert np.abs(got32['v_x'] - want['v_x']).max() < 1e-2 and np.abs(got32['v_y'] - want['v_y']).max() < 1e-2
    assert got64['solver_info']['iterations'][0] <= 40


@pytest.mark.parametrize('variant', ['default', 'cells_split', 'batch_spans_pairs', 'two_threads', 'many_pairs'])
def test_vary_regularisation_matches_loop(cuda_device, variant):
    """2 pairs x 6 cells against a loop of oracle calls (optical_flow.py:1960-1983): the cells of a pair in one batch,
    a pair's cells split over batches, one batch spanning both pairs, the pairs sharded over two host threads; and 11 pairs
    in batches of 8 pairs of one cell each (the layout long sweeps use), the last batch a short one."""
    movie = synthetic.translating_movie(40, 48, 12 if variant == 'many_pairs' else 3, seed=12)
    alphas, betas = np.array([2e4, 1e5]), np.array([1e3, 1e4, 1e5])
    extra = {'default': {}, 'cells_split': dict(batch_pairs=4), 'batch_spans_pairs': dict(batch_pairs=12),
             'two_threads': dict(devices=['cuda:0', 'cuda:0']), 'many_pairs': dict(batch_pairs=8)}[variant]
    got = optical_flow.vary_regularisation(movie, alphas, betas, smoothing_sigma=2.0, rtol=1e-11, max_it=20000, **extra)
    assert got['speed_means'].shape == (2, 3) and got['converged'].dtype == bool
    for i, a in enumerate(alphas):
        for j, b in enumerate(betas):
            ref = fo.variational_optical_flow(movie, speed_alpha=a, remodelling_alpha=b, smoothing_sigma=2.0)
            assert abs(got['speed_means'][i, j] - np.mean(ref['speed'])) < 1e-7 * np.mean(ref['speed'])
            assert abs(got['speed_variances'][i, j] - np.var(ref['speed'])) < 1e-6 * np.var(ref['speed'])
            assert abs(got['remodelling_means'][i, j] - np.mean(ref['remodelling'])) < 1e-6 * abs(np.mean(ref['remodelling'])) + 1e-12
            total = ref['L1_functional'] + ref['speed_functional'] + ref['remodelling_functional']
            assert abs(got['functional'][i, j] - total) < 1e-6 * total
            assert got['converged'][i, j]


def test_vary_regularisation_128_two_pairs_six_cells_float32(cuda_device):
    """The sweep at 128^2 in float32 (what config d-4 runs), 2 pairs x 6 cells in one 12-entry solve, against the oracle loop."""
    movie = synthetic.rotating_shearing_movie(128, 128, 3, seed=2)
    alphas, betas = np.array([4.64e4, 1e5, 1e6]), np.array([1e3, 1e4])
    got = optical_flow.vary_regularisation(movie, alphas, betas, smoothing_sigma=2.5, dtype='float32', rtol=1e-6)
    for i, a in enumerate(alphas):
        for j, b in enumerate(betas):
            ref = fo.variational_optical_flow(movie, speed_alpha=a, remodelling_alpha=b, smoothing_sigma=2.5)
            assert abs(got['speed_means'][i, j] - np.mean(ref['speed'])) < 1e-3 * np.mean(ref['speed'])
            assert abs(got['remodelling_means'][i, j] - np.mean(ref['remodelling'])) < 1e-3 * abs(np.mean(ref['remodelling'])) + 1e-6
            total = ref['L1_functional'] + ref['speed_functional'] + ref['remodelling_functional']
            assert abs(got['functional'][i, j] - total) < 2e-3 * total
            assert got['converged'][i, j]


def test_float32_pairs_that_stall_are_redone_in_float64(cuda_device):
    """alpha close to max(I^2): the float32 solve stalls above rtol, the drop-in redoes the pair in float64
    (solver_info['fallback']) and the result still meets the float32 tolerance against the oracle."""
    movie = synthetic.rotating_shearing_movie(128, 128, 2, seed=2)
    kw = dict(speed_alpha=2.15e4, remodelling_alpha=1e4, smoothing_sigma=2.5)
    want = fo.variational_optical_flow(movie, **kw)
    got = optical_flow.variational_optical_flow(movie, dtype='float32', rtol=1e-6, max_it=400, **kw)
    assert got['converged']
    for key in ('v_x', 'v_y'):
        assert rel(got[key], want[key]) < 1e-3
        assert np.abs(got[key] - want[key]).max() < 1e-2


@pytest.mark.parametrize('N,alpha', [(256, 1e3), (128, 4.64e3), (128, 2.15e4)])
def test_low_alpha_regime_matches_the_oracle(cuda_device, N, alpha):
    """speed_alpha below ~max I^2 / 1.5: the operator is indefinite (profiles/r02_indefinite_regime.txt), the multigrid
    BiCGStab does not converge, and the pairs go to the sparse direct solve - no rescue in the test.  The reference's real-data
    regime (alpha = 2e4 on 8-bit frames, analysis/analyse_variational_optical_flow.py:114) and config d-4's lowest cells."""
    movie = synthetic.rotating_shearing_movie(N, N, 2, seed=2)
    kw = dict(speed_alpha=alpha, remodelling_alpha=1e4, smoothing_sigma=2.5)
    want = fo.variational_optical_flow(movie, **kw)
    got64 = optical_flow.variational_optical_flow(movie, dtype='float64', rtol=1e-10, **kw)         # parity setting of north_star
    got32 = optical_flow.variational_optical_flow(movie, dtype='float32', rtol=1e-6, **kw)
    for got, tol in ((got64, 1e-6), (got32, 1e-3)):
        assert got['converged'] and got['solver_info']['converged'].all()
        for key in ('v_x', 'v_y', 'remodelling'):
            assert rel(got[key], want[key]) < tol, (key, got['solver_info'])
    assert np.abs(got32['v_x'] - want['v_x']).max() < 1e-2 and np.abs(got32['v_y'] - want['v_y']).max() < 1e-2
    if N == 256:
        assert got64['solver_info']['fallback'].all()          # 256^2 at alpha = 1e3: the Krylov path cannot have converged


@pytest.mark.parametrize('name', GOLDEN_NAMES)
def test_use_direct_solver_reproduces_the_reference_direct_branch(cuda_device, name):
    """use_direct_solver=True (:1146-1147): the device's sparse direct solve against the golden outputs of the reference's own
    SuperLU branch - no iteration, no tolerance to choose."""
    z, kwargs = load_golden(name)
    got = optical_flow.variational_optical_flow(z['movie'], use_direct_solver=True, **kwargs)
    assert got['solver_info']['direct_solver'] and got['converged'] and (got['solver_info']['iterations'] == 0).all()
    for key in ('v_x', 'v_y', 'speed', 'remodelling'):
        assert rel(got[key], z[key]) < 1e-8, key
    fun = np.array([got['L1_functional'], got['remodelling_functional'], got['speed_functional']])
    np.testing.assert_allclose(fun, z['functionals'], rtol=1e-8)
    assert got['solver_info']['relative_residual'].max() < 1e-9


# ------------------------------------------------------------------ bench sizes, anchored to the ORACLE operator
# The reference's own check of a solve is  |LHS_matrix @ x - RHS| / |RHS|  (optical_flow.py:1150-1151).  The oracle's
# apply_operator / build_rhs are that matrix and right-hand side bit for bit (tests/test_oracle_cpu.py), and cost
# seconds of numpy at 1024^2 - no solve - so every size the bench runs is held to them here.
U32, U64 = 2.0 ** -24, 2.0 ** -53


def raw_solution(result, k, scale=1.0):
    """The linear system's solution behind the returned fields of pair k: apply_constant_boundary_condition (:1164-1166)
    only rewrites the ghost ring, and the ring of the raw solution follows from the boundary rows (:965-1070):
    q(0,j) = q(2,j), ..., corners q(0,0) = q(2,0) + q(0,2)."""
    x = np.stack([result['v_x'][k] / scale, result['v_y'][k] / scale, result['remodelling'][k]]).astype(np.float64)
    return mg.fill_ghosts(x)


def oracle_residual(P, C, x, alpha, beta):
    """(|b - A x| / |b|,  u-free rounding scale |D x| / |b|) with the oracle's matrix-free restatement of :836-1072."""
    b = fo.build_rhs(P, C)
    r = b - fo.apply_operator(P, x, alpha, beta)
    I, Ix, Iy, Ixx, Iyy, Ixy = fo.image_coefficients(P)
    inner = (slice(1, -1), slice(1, -1))
    dx = np.stack([(I * (Ixx - 2 * I) - 4 * alpha) * x[0][inner], (I * (Iyy - 2 * I) - 4 * alpha) * x[1][inner],
                   (-1 - 4 * beta) * x[2][inner]])
    bn = np.linalg.norm(b)
    return np.linalg.norm(r) / bn, np.linalg.norm(dx) / bn


def check_movie_against_oracle_operator(movie, kw, iteration_cap):
    """fp64 at rtol 1e-10 and fp32 at rtol 1e-6 through the drop-in; residuals through the oracle operator."""
    alpha, beta, sigma = kw['speed_alpha'], kw['remodelling_alpha'], kw['smoothing_sigma']
    got64 = optical_flow.variational_optical_flow(movie, dtype='float64', rtol=1e-10, **kw)
    got32 = optical_flow.variational_optical_flow(movie, dtype='float32', rtol=1e-6, **kw)
    assert got64['solver_info']['converged'].all() and got32['solver_info']['converged'].all()
    assert not got64['solver_info']['fallback'].any() and not got32['solver_info']['fallback'].any()
    assert got32['solver_info']['iterations'].max() <= iteration_cap
    blurred = fo.blur_movie(movie, sigma)                               # the oracle's own blur, float64
    assert np.abs(got64['blurred_data'] - blurred).max() < 1e-9
    for k in range(movie.shape[0] - 1):
        # float64: everything on the oracle's side is the oracle's
        res, _ = oracle_residual(blurred[k], blurred[k + 1], raw_solution(got64, k), alpha, beta)
        assert res < 1e-9, (k, res)
        assert abs(got64['solver_info']['relative_residual'][k] - res) < 1e-2 * res + 1e-13
        # float32: the frames the device solved on are its own float32 blur (returned widened, exactly); x is stored in
        # float32, so |b - A x| cannot go below ~u |D x| whatever the solver does - the bar is the backward-error form
        # rtol |b| + 8 u |D x| (include/ofb200.h), and the forward error against the float64 answer the stated 1e-3 / 1e-2 px
        P32, C32 = got32['blurred_data'][k], got32['blurred_data'][k + 1]
        res32, dscale = oracle_residual(P32, C32, raw_solution(got32, k), alpha, beta)
        assert res32 <= 1e-6 + 8 * U32 * dscale, (k, res32, U32 * dscale)
        assert abs(got32['solver_info']['relative_residual'][k] - res32) < 1e-3 * res32
        assert got32['solver_info']['recurrence_residual'][k] <= 1e-6
    for key in ('v_x', 'v_y', 'remodelling'):
        assert rel(got32[key], got64[key]) < 1e-3, key
    assert np.abs(got32['v_x'] - got64['v_x']).max() < 1e-2 and np.abs(got32['v_y'] - got64['v_y']).max() < 1e-2
    return got64, got32


def test_config_d3_1024_residual_through_oracle_operator(cuda_device):
    """BASELINE config d-3 (the headline workload): 1024^2 rotating + shearing movie."""
    movie = synthetic.rotating_shearing_movie(1024, 1024, 3)
    got64, _ = check_movie_against_oracle_operator(movie, dict(speed_alpha=1e5, remodelling_alpha=1e4, smoothing_sigma=2.5), 30)
    # recovered flow is close to the imposed rotation+shear in the interior
    ii, jj = np.meshgrid(np.arange(1024.0), np.arange(1024.0), indexing='ij')
    vx_true = (-1e-3 + 5e-4) * (jj - 511.5)
    sl = (slice(100, -100), slice(100, -100))
    assert np.abs(got64['v_x'][0][sl] - vx_true[sl]).mean() < 0.1


def test_config_d2_512_actin_against_oracle_operator_and_reference_sample(cuda_device):
    """BASELINE config d-2: 512^2 actin-like movie.  Residuals through the oracle operator, and pair 0 against a sample of
    the UNMODIFIED reference's own SuperLU answer (tests/golden/actin512_pair0_sample.npz, oracle/make_golden.py
    --large-only: every 8th pixel, the central 64 x 64 block, the three outermost rows and columns)."""
    import os
    from conftest import GOLDEN_DIR
    movie = synthetic.actin_like_movie(512, 512, 3, seed=1)
    kw = dict(speed_alpha=1e5, remodelling_alpha=1e4, smoothing_sigma=2.5)
    got64, got32 = check_movie_against_oracle_operator(movie, kw, 30)
    z = np.load(os.path.join(GOLDEN_DIR, 'actin512_pair0_sample.npz'))
    np.testing.assert_allclose([movie[:2].sum(), (movie[:2] ** 2).sum()], z['movie_checksum'], rtol=1e-12)
    for got, tol in ((got64, 1e-6), (got32, 1e-3)):
        for key in ('v_x', 'v_y', 'speed', 'remodelling'):
            f = got[key][0]
            views = dict(sub8=f[::8, ::8], centre=f[224:288, 224:288], top=f[:3], bottom=f[-3:], left=f[:, :3], right=f[:, -3:])
            for name, view in views.items():
                # every sample is judged on the scale of the whole field (a near-zero patch has no relative accuracy)
                err = np.linalg.norm(view - z[key + '_' + name]) / (float(z[key + '_norm']) * np.sqrt(view.size / f.size))
                assert err < tol, (key, name, err)


@pytest.mark.parametrize('shape', [(1024, 1024), (16, 8192), (300, 2052)])
@pytest.mark.parametrize('dtype,tol', [(torch.float64, 1e-13), (torch.float32, 2e-5)])
def test_full_operator_matches_oracle_at_bench_sizes(cuda_device, shape, dtype, tol):
    """ofb_apply and ofb_apply_eliminated at 1024^2 and at the 8192-column width of config d-5."""
    rng = np.random.default_rng(shape[1])
    P = fo.gaussian_blur(rng.uniform(20, 200, shape), 1.5)
    x = rng.standard_normal((3,) + shape)
    alpha, beta = 1e5, 1e4
    Pd, xd = dev_t(P[None], cuda_device, dtype), dev_t(x[None], cuda_device, dtype)
    got = engine.apply_operator(Pd, xd, alpha, beta)[0].cpu().double().numpy()
    assert rel(got, fo.apply_operator(P, x, alpha, beta)) < tol
    got_e = engine.apply_operator(Pd, xd, alpha, beta, eliminated=True)[0].cpu().double().numpy()
    want_e = fo.apply_operator(P, mg.fill_ghosts(x.copy()), alpha, beta)      # ring substituted: boundary rows vanish
    assert rel(got_e[:, 1:-1, 1:-1], want_e[:, 1:-1, 1:-1]) < tol and np.abs(got_e[:, 0, :]).max() == 0.0


@pytest.mark.parametrize('dtype', [torch.float64, torch.float32])
def test_true_residual_output_is_the_oracle_residual(cuda_device, dtype):
    """ofb_solve's relres_out is |b - A x| / |b| of the x it returns (reference :1150-1151), recomputed in fp64
    arithmetic whatever the storage type; recurrence_out is the solver's own norm (:1140-1142)."""
    movie = fo.blur_movie(synthetic.translating_movie(96, 132, 3, seed=21), 2.5)
    alphas, betas = np.array([1e5, 3e4]), np.array([1e4, 1e3])
    Pd, Cd = dev_t(movie[:-1], cuda_device, dtype), dev_t(movie[1:], cuda_device, dtype)
    b = engine.build_rhs(Pd, Cd)
    x = torch.zeros_like(b)
    rec = torch.zeros(2, dtype=torch.float64, device=cuda_device)
    rtol = 1e-10 if dtype == torch.float64 else 1e-6
    it, rr, cv = engine.solve(Pd, b, x, torch.from_numpy(alphas), torch.from_numpy(betas), rtol=rtol, max_it=500, recurrence=rec)
    assert cv.tolist() == [1, 1]
    for k in range(2):
        P, C = Pd[k].cpu().double().numpy(), Cd[k].cpu().double().numpy()
        res, dscale = oracle_residual(P, C, x[k].cpu().double().numpy(), alphas[k], betas[k])
        if dtype == torch.float32:          # b was built in float32 on the device: compare on the device's own b
            bk = b[k].cpu().double().numpy()
            res = np.linalg.norm(bk - fo.apply_operator(P, x[k].cpu().double().numpy(), alphas[k], betas[k])) / np.linalg.norm(bk)
        assert abs(float(rr[k]) - res) < 1e-3 * res + 1e-14, (float(rr[k]), res)
        assert float(rec[k]) <= rtol
        assert res <= rtol + 8 * (U64 if dtype == torch.float64 else U32) * dscale


def test_identical_frames_give_zero_whatever_the_guess(cuda_device):
    """b = 0 (a duplicated frame): the solution is x = 0; PETSc's default test would iterate from the guess towards it
    (:1092, :1120), returning the guess as 'converged' would not."""
    frame = synthetic.textured_frame(40, 48, seed=4)
    movie = np.stack([frame, frame, synthetic.fourier_shift(frame, 0.5, 0.3)])
    got = optical_flow.variational_optical_flow(movie, speed_alpha=1e5, remodelling_alpha=1e4, initial_v_x=0.7,
                                                initial_v_y=-0.2, initial_remodelling=0.1, rtol=1e-10)
    assert got['solver_info']['converged'].all()
    for key in ('v_x', 'v_y', 'speed', 'remodelling'):
        # the reference's own difference order (:815-819) leaves round-off of ~1e-14 in b: a solve, and a solution of ~1e-18
        assert np.abs(got[key][0]).max() < 1e-12
    want = fo.variational_optical_flow(movie[1:], speed_alpha=1e5, remodelling_alpha=1e4)
    assert rel(got['v_x'][1], want['v_x'][0]) < 1e-6
    # b exactly 0 through the C ABI: x = 0, reported converged, no iteration
    P = dev_t(movie[:1], cuda_device)
    b = torch.zeros((1, 3, 40, 48), dtype=torch.float64, device=cuda_device)
    x = torch.full_like(b, 0.7)
    it, rr, cv = engine.solve(P, b, x, 1e5, 1e4, rtol=1e-10, max_it=50)
    assert cv.tolist() == [1] and it.tolist() == [0] and float(x.abs().max()) == 0.0


def test_devices_argument_on_two_distinct_gpus(cuda_device):
    """devices=['cuda:0', 'cuda:1']: one host thread per GPU, each making its GPU current (the library keys its per-device
    state on cudaGetDevice()).  Needs two GPUs."""
    if torch.cuda.device_count() < 2:
        pytest.skip('needs two GPUs')
    movie = synthetic.translating_movie(96, 128, 7, seed=9)
    kw = dict(speed_alpha=1e5, remodelling_alpha=1e4, smoothing_sigma=2.0, rtol=1e-10)
    one = optical_flow.variational_optical_flow(movie, device='cuda:1', **kw)
    two = optical_flow.variational_optical_flow(movie, devices=['cuda:0', 'cuda:1'], **kw)
    for key in ('v_x', 'v_y', 'speed', 'remodelling'):
        assert rel(two[key], one[key]) < 1e-7, key


# ------------------------------------------------------------------ streamed (TMA bulk copy + mbarrier ring) sweeps
# shapes wide enough for the streamed path (Ny >= 64 vector lanes, Ny % 4 == 0), with ragged tile ends,
# single-column last tiles, several strips and strips of one row
WIDE_SHAPES = [(9, 256), (70, 260), (33, 512), (130, 516), (64, 1024), (21, 1028), (257, 772)]


@pytest.mark.parametrize('shape', WIDE_SHAPES)
@pytest.mark.parametrize('dtype,tol', [(torch.float64, 1e-12), (torch.float32, 3e-5)])
def test_streamed_smoother_matches_model(cuda_device, shape, dtype, tol):
    """x_out = x + w D^-1 (b - S x) through the streamed sweep == numpy model; and == the register sweep."""
    from opticalflow_b200 import _lib
    rng = np.random.default_rng(shape[0] * 31 + shape[1])
    B = 3
    P = fo.blur_movie(rng.uniform(20, 200, (B,) + shape), 1.5)          # smooth image: well-conditioned 3x3 blocks
    x = mg.zero_ghosts(rng.standard_normal((B, 3) + shape))
    b = mg.zero_ghosts(rng.standard_normal((B, 3) + shape) * 1e4)
    alphas, betas = np.array([1e5, 1e3, 2e4]), np.array([1e4, 1e2, 1e3])
    Pd, xd, bd = dev_t(P, cuda_device, dtype), dev_t(x, cuda_device, dtype), dev_t(b, cuda_device, dtype)
    at, bt = torch.from_numpy(alphas), torch.from_numpy(betas)
    for rows in (0, 1, 7):
        got = engine.smooth(Pd, bd, xd, at, bt, omega=0.8, rows_per_strip=rows).cpu().double().numpy()
        reg = engine.smooth(Pd, bd, xd, at, bt, omega=0.8, rows_per_strip=rows, flags=_lib.FLAG_NO_STREAM).cpu().double().numpy()
        for k in range(B):
            r = b[k] - mg.apply_eliminated(P[k], x[k], alphas[k], betas[k])
            want = mg.zero_ghosts(x[k] + 0.8 * mg.apply_block_jacobi(P[k], r, alphas[k], betas[k]))
            assert rel(got[k], want) < tol, (rows, k)
            assert rel(got[k], reg[k]) < tol


@pytest.mark.parametrize('shape', [(33, 512), (130, 516), (64, 1024)])
@pytest.mark.parametrize('dtype,tol', [(torch.float64, 1e-13), (torch.float32, 2e-5)])
def test_streamed_operator_matches_model(cuda_device, shape, dtype, tol):
    rng = np.random.default_rng(shape[0] * 13 + shape[1])
    P = rng.uniform(20, 200, (2,) + shape)
    x = rng.standard_normal((2, 3) + shape)
    got = engine.apply_operator(dev_t(P, cuda_device, dtype), dev_t(x, cuda_device, dtype), 1e5, 1e4, eliminated=True)
    got = got.cpu().double().numpy()
    for k in range(2):
        assert rel(got[k], mg.apply_eliminated(P[k], x[k], 1e5, 1e4)) < tol


@pytest.mark.parametrize('precond', ['mg', 'bjacobi'])
def test_streamed_solve_matches_register_solve_and_direct(cuda_device, precond):
    """A wide ragged frame (streamed levels 0 and 1, register sweeps below) against the direct solution."""
    from opticalflow_b200 import _lib
    shape = (48, 516)
    movie = fo.blur_movie(synthetic.translating_movie(shape[0], shape[1], 2, seed=21), 2.5)
    P, C = movie
    xd, _ = fo.solve_pair(P, C, 1e5, 1e4)
    Pd, Cd = dev_t(P[None], cuda_device), dev_t(C[None], cuda_device)
    b = engine.build_rhs(Pd, Cd)
    for flags in (0, _lib.FLAG_NO_STREAM):
        x = torch.zeros_like(b)
        it, rr, cv = engine.solve(Pd, b, x, 1e5, 1e4, rtol=1e-11, max_it=20000, options=engine.make_options(precond, flags=flags))
        assert int(cv[0]) == 1, (flags, int(it[0]), float(rr[0]))
        assert rel(x[0].cpu().numpy(), xd) < 1e-7, flags


@pytest.mark.parametrize('shape', [(66, 66), (130, 98), (258, 516)])
@pytest.mark.parametrize('dtype,tol', [(torch.float64, 1e-11), (torch.float32, 2e-4)])
def test_fused_coarse_tail_matches_per_level_kernels(cuda_device, shape, dtype, tol):
    """One V-cycle application with the shared-memory coarse tail == the same cycle run level by level."""
    from opticalflow_b200 import _lib
    movie = fo.blur_movie(synthetic.translating_movie(shape[0], shape[1], 3, seed=31), 2.0)
    rng = np.random.default_rng(4)
    r = mg.zero_ghosts(rng.standard_normal((2, 3) + shape))
    Pd, rd = dev_t(movie[:2], cuda_device, dtype), dev_t(r, cuda_device, dtype)
    alphas, betas = torch.tensor([1e5, 2e4], dtype=torch.float64), torch.tensor([1e4, 1e3], dtype=torch.float64)
    for cycle in (dict(), dict(nu_pre=2, nu_post=2, omega=0.8)):
        z_fused = engine.precondition(Pd, rd, alphas, betas, options=engine.make_options('mg', **cycle))
        z_plain = engine.precondition(Pd, rd, alphas, betas,
                                      options=engine.make_options('mg', flags=_lib.FLAG_NO_FUSED_TAIL, **cycle))
        assert rel(z_fused.cpu().double().numpy(), z_plain.cpu().double().numpy()) < tol, cycle


def test_devices_argument_shards_pairs_over_threads(cuda_device):
    """devices=[...] (one host thread per device; here twice the same GPU) == the unsharded call."""
    movie = synthetic.translating_movie(48, 64, 6, seed=17)
    kw = dict(speed_alpha=1e5, remodelling_alpha=1e4, smoothing_sigma=2.0, dtype='float64', rtol=1e-11, max_it=2000)
    one = optical_flow.variational_optical_flow(movie, **kw)
    two = optical_flow.variational_optical_flow(movie, devices=['cuda:0', 'cuda:0'], **kw)
    for key in ('v_x', 'v_y', 'speed', 'remodelling', 'blurred_data', 'original_data'):
        assert two[key].shape == one[key].shape
        assert rel(two[key], one[key]) < 1e-9, key
    assert two['solver_info']['shards'] == [3, 2] and two['converged']
    assert abs(two['L1_functional'] - one['L1_functional']) < 1e-8 * abs(one['L1_functional'])


def test_direct_and_staged_egress_return_identical_arrays(cuda_device, monkeypatch):
    """Direct egress (page-locked result arrays, float64 widening on the device) against the staged path (pinned float32
    staging, widening by the host copy pool), for a float32 and a float64 solve, over several chunks."""
    movie = synthetic.translating_movie(96, 128, 6, seed=5).astype(np.float32)
    kw = dict(speed_alpha=1e5, remodelling_alpha=1e4, smoothing_sigma=2.0, batch_pairs=2)
    for dtype in ('float32', 'float64'):
        monkeypatch.setenv('OFB_DIRECT_EGRESS', '1')
        direct = optical_flow.variational_optical_flow(movie, dtype=dtype, **kw)
        assert direct['solver_info']['direct_egress']
        monkeypatch.setenv('OFB_DIRECT_EGRESS', '0')
        staged = optical_flow.variational_optical_flow(movie, dtype=dtype, **kw)
        assert not staged['solver_info']['direct_egress']
        # the blur has no atomics: bit-identical planes through both paths; two solves of the same pair differ in the
        # last bits (dot products are summed by atomics), so the fields agree to the solver tolerance, not to the bit
        for key in ('blurred_data', 'original_data'):
            assert direct[key].dtype == np.float64 and np.array_equal(direct[key], staged[key]), key
        for key in ('v_x', 'v_y', 'speed', 'remodelling'):
            assert direct[key].dtype == np.float64 and direct[key].shape == staged[key].shape
            assert rel(direct[key], staged[key]) < (1e-4 if dtype == 'float32' else 1e-6), key
        if dtype == 'float32':      # widened on the device or on the host, every value is a float32
            assert np.array_equal(direct['v_x'], direct['v_x'].astype(np.float32).astype(np.float64))
        assert np.array_equal(direct['original_data'], movie.astype(np.float64))
    # dropped results hand their page-locked buffers back to the arena; the next call of the same shape reuses them
    import gc
    from opticalflow_b200 import hostio
    monkeypatch.setenv('OFB_DIRECT_EGRESS', '1')
    del direct, staged
    gc.collect()
    idle = len(hostio.result_arena.free_pinned)
    again = optical_flow.variational_optical_flow(movie, dtype='float32', **kw)
    assert again['solver_info']['direct_egress'] and idle >= 1 and len(hostio.result_arena.free_pinned) < idle


def test_warm_start_changes_iterations_not_answers(cuda_device):
    """warm_start=True (the reference's seeding of a pair with an earlier solution, :803-806, across chunks): same fields
    within the solver tolerance, fewer iterations on a stationary flow."""
    movie = synthetic.translating_movie(96, 128, 9, seed=5)
    kw = dict(speed_alpha=1e5, remodelling_alpha=1e4, smoothing_sigma=2.0, batch_pairs=2, rtol=1e-9)
    cold = optical_flow.variational_optical_flow(movie, **kw)
    warm = optical_flow.variational_optical_flow(movie, warm_start=True, **kw)
    assert cold['solver_info']['converged'].all() and warm['solver_info']['converged'].all()
    for key in ('v_x', 'v_y', 'remodelling'):
        assert rel(warm[key], cold[key]) < 1e-6, key
    assert (warm['solver_info']['iterations'][:2] == cold['solver_info']['iterations'][:2]).all()       # first chunk: same start
    assert warm['solver_info']['iterations'][2:].sum() < cold['solver_info']['iterations'][2:].sum()


def test_result_dtype_and_outputs_keep_the_values(cuda_device, monkeypatch):
    """Opt-in contracts: result_dtype='float32' and outputs=(...) return the same numbers as the reference-shaped default
    (to float32 rounding), leave out exactly the keys not asked for, on the direct and on the staged egress path; and
    speed computed on the host (OFB_HOST_SPEED=1) is sqrt(v_x^2 + v_y^2) of the returned arrays."""
    movie = synthetic.translating_movie(96, 128, 6, seed=5).astype(np.float32)
    kw = dict(speed_alpha=1e5, remodelling_alpha=1e4, smoothing_sigma=2.0, batch_pairs=2, dtype='float32')
    full = optical_flow.variational_optical_flow(movie, **kw)
    for direct in ('1', '0'):
        monkeypatch.setenv('OFB_DIRECT_EGRESS', direct)
        lean = optical_flow.variational_optical_flow(movie, result_dtype='float32', outputs=('v_x', 'v_y', 'remodelling'), **kw)
        assert lean['solver_info']['direct_egress'] == (direct == '1')
        assert not ({'speed', 'blurred_data', 'original_data'} & set(lean))
        for key in ('v_x', 'v_y', 'remodelling'):
            assert lean[key].dtype == np.float32 and rel(lean[key], full[key]) < 1e-4, key
        assert lean['converged'] == full['converged'] and abs(lean['L1_functional'] - full['L1_functional']) < 1e-3 * full['L1_functional']
        assert lean['solver_info']['d2h_bytes'] * 42 < full['solver_info']['d2h_bytes'] * 13       # 12 of the 40 bytes per pixel
        f32 = optical_flow.variational_optical_flow(movie, result_dtype='float32', **kw)
        assert set(f32) == set(full) and f32['original_data'] is not None
        for key in ('v_x', 'v_y', 'speed', 'remodelling', 'blurred_data', 'original_data'):
            assert f32[key].dtype == np.float32 and rel(f32[key], full[key]) < 1e-4, key
    monkeypatch.setenv('OFB_DIRECT_EGRESS', '1')
    monkeypatch.setenv('OFB_HOST_SPEED', '1')
    host = optical_flow.variational_optical_flow(movie, **kw)
    assert host['solver_info']['speed_on_host'] and host['speed'].dtype == np.float64
    assert np.abs(host['speed'] - np.sqrt(host['v_x'] ** 2 + host['v_y'] ** 2)).max() < 1e-14
    assert rel(host['speed'], full['speed']) < 1e-4
    with pytest.raises(ValueError):
        optical_flow.variational_optical_flow(movie, outputs=('v_x', 'vorticity'), **kw)


@pytest.mark.parametrize('shape', [(64, 256), (130, 512), (257, 1028), (96, 1024)])
def test_fused_prolongation_matches_separate_pass(cuda_device, shape):
    """float32 image level, OFB_FLAG_FUSED_PROLONG: the first post-smoothing sweep adds the coarse-grid correction to its
    window as it loads it (same arithmetic as the separate pass, one launch fewer per V-cycle).  One preconditioner
    application and a whole solve against the default path."""
    from opticalflow_b200 import _lib
    rng = np.random.default_rng(shape[1])
    B = 3
    P = dev_t(rng.uniform(20, 200, (B,) + shape), cuda_device, torch.float32)
    r = dev_t(rng.standard_normal((B, 3) + shape), cuda_device, torch.float32)
    r[:, :, 0, :] = 0; r[:, :, -1, :] = 0; r[:, :, :, 0] = 0; r[:, :, :, -1] = 0
    alpha, beta = torch.tensor([1e5, 3e5, 1e6]), torch.tensor([1e4, 1e3, 1e5])
    lib = _lib.load()
    n0 = lib.ofb_launch_count()
    plain = engine.precondition(P, r, alpha, beta, engine.make_options('mg'))
    n1 = lib.ofb_launch_count()
    fused = engine.precondition(P, r, alpha, beta, engine.make_options('mg', flags=_lib.FLAG_FUSED_PROLONG))
    n2 = lib.ofb_launch_count()
    assert (n2 - n1) == (n1 - n0) - 1                                    # the fused path ran
    inner = (slice(None), slice(None), slice(1, -1), slice(1, -1))
    assert float((fused[inner] - plain[inner]).abs().max()) <= 1e-5 * float(plain[inner].abs().max())
    xs = []
    for flags in (0, _lib.FLAG_FUSED_PROLONG):
        x = torch.zeros_like(r)
        it, rr, cv = engine.solve(P, r, x, alpha, beta, 1e-6, 200, engine.make_options('mg', flags=flags))
        assert int(cv.sum()) == B
        xs.append((x.cpu().double().numpy(), it.cpu().numpy()))
    assert rel(xs[0][0], xs[1][0]) < 1e-4 and np.abs(xs[0][1] - xs[1][1]).max() <= 1
