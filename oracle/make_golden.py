"""TEST INFRASTRUCTURE - generates tests/golden/*.npz by running the UNMODIFIED reference
(/root/reference/source/optical_flow.py behind stubs, oracle/ref_import.py) through its own
``use_direct_solver=True`` branch.  Run in the build container only:  python oracle/make_golden.py

Each fixture stores the inputs needed to regenerate the movie (generator name + arguments), the kwargs and the
reference outputs (v_x, v_y, speed, remodelling, functionals, blurred_data for the blurred cases).
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_import  # noqa: E402
from opticalflow_b200 import synthetic  # noqa: E402

GOLDEN = os.path.join(ROOT, 'tests', 'golden')

CASES = {
    # analysis/analyse_variational_optical_flow.py:26-50 (simple_test_with_data_on_boundary)
    'hat50': dict(gen='hat', gen_args=dict(dimension=50),
                  kwargs=dict(delta_t=1.0, speed_alpha=1.0, remodelling_alpha=10000.0, smoothing_sigma=None)),
    # SURVEY d-1 family at reduced size
    'texture64': dict(gen='translate', gen_args=dict(Nx=64, Ny=64, n_frames=2),
                      kwargs=dict(speed_alpha=1e5, remodelling_alpha=1e4, smoothing_sigma=2.5)),
    # ragged, non-square, three frames (two pairs), no blur, delta_x/delta_t scaling
    'ragged_40x56': dict(gen='translate', gen_args=dict(Nx=40, Ny=56, n_frames=3, seed=7),
                         kwargs=dict(delta_x=0.5, delta_t=2.0, speed_alpha=2e4, remodelling_alpha=1e3,
                                     smoothing_sigma=None)),
    # real-data regime of analyse_variational_optical_flow.py:114,:139-145 (sigma = 0.62*4)
    'texture96_realregime': dict(gen='translate', gen_args=dict(Nx=96, Ny=80, n_frames=2, seed=3),
                                 kwargs=dict(speed_alpha=1e5, remodelling_alpha=1e4, smoothing_sigma=0.62 * 4)),
    # smallest frame the path accepts
    'tiny_5x7': dict(gen='translate', gen_args=dict(Nx=5, Ny=7, n_frames=2, seed=11),
                     kwargs=dict(speed_alpha=1e3, remodelling_alpha=1e2, smoothing_sigma=None)),
}


def make_movie(case):
    if case['gen'] == 'hat':
        movie, delta_x = synthetic.gaussian_hat_pair(**case['gen_args'])
        return movie, dict(delta_x=float(delta_x))
    return synthetic.translating_movie(**case['gen_args']), {}


def main():
    os.makedirs(GOLDEN, exist_ok=True)
    for name, case in CASES.items():
        movie, extra = make_movie(case)
        kwargs = dict(case['kwargs'], **extra)
        result, residuals, _ = ref_import.reference_solve(movie, **kwargs)
        payload = dict(movie=movie.astype(np.float64), residuals=np.asarray(residuals))
        for key in ('v_x', 'v_y', 'speed', 'remodelling'):
            payload[key] = result[key]
        if kwargs.get('smoothing_sigma') is not None:
            payload['blurred_data'] = result['blurred_data']
        payload['functionals'] = np.array([result['L1_functional'], result['remodelling_functional'],
                                           result['speed_functional']])
        for k, v in kwargs.items():
            payload['kw_' + k] = np.array(np.nan if v is None else v, dtype=np.float64)
        np.savez_compressed(os.path.join(GOLDEN, name + '.npz'), **payload)
        print(name, movie.shape, 'max residual %.2e' % max(residuals))


def _texture(T, Nx, Ny, seed, lo=20.0, hi=220.0):
    """Smooth random texture drifting by one pixel per frame plus a slow brightening (stored in the fixture)."""
    import scipy.ndimage
    rng = np.random.default_rng(seed)
    base = scipy.ndimage.gaussian_filter(rng.standard_normal((Nx + 2 * T, Ny + 2 * T)), 2.0)
    base = (base - base.min()) / (base.max() - base.min()) * (hi - lo) + lo
    return np.stack([base[T + t:T + t + Nx, T:T + Ny] + 0.3 * t for t in range(T)])


def make_legacy():
    """Fixtures for the callers either side of the hot path (SURVEY 8f NEXT-3/NEXT-4): the reference's own numba
    (liu_shen_optical_flow_jit) and OpenCV (apply_adaptive_threshold, apply_clahe) code, run unmodified here."""
    import contextlib
    import io
    ref = ref_import.import_reference()
    rng = np.random.default_rng(42)
    payload = {}
    # --- Liu-Shen: ragged three-frame movie with random guesses and unit scaling, and the driver's constant guesses
    movie = _texture(3, 22, 19, 1)
    gx, gy, gr = (rng.standard_normal((22, 19)) * 0.1 for _ in range(3))
    out = ref.liu_shen_optical_flow_jit(movie, 0.5, 2.0, 100.0, 1.0, gx, gy, gr, 7, 1e-9, True)
    payload.update(ls_a_movie=movie, ls_a_gx=gx, ls_a_gy=gy, ls_a_gr=gr, ls_a_args=np.array([0.5, 2.0, 100.0, 7]),
                   ls_a_vx=out[0], ls_a_vy=out[1], ls_a_speed=out[2], ls_a_rem=out[3], ls_a_last=np.array(out[4]))
    movie = _texture(2, 16, 136, 2)          # wider than one 128-column tile
    g = np.full((16, 136), 0.1)
    out = ref.liu_shen_optical_flow_jit(movie, 1.0, 1.0, 1e4, 1.0, g, g, np.full((16, 136), 0.5), 12, 1e-9, True)
    payload.update(ls_b_movie=movie, ls_b_args=np.array([1.0, 1.0, 1e4, 12]), ls_b_vx=out[0], ls_b_vy=out[1], ls_b_speed=out[2])
    # --- adaptive threshold (defaults; and the analysis scripts' window 151 / threshold -5)
    movie = np.abs(_texture(2, 70, 90, 3)) ** 1.5
    payload.update(th_a_movie=movie, th_a_args=np.array([51, 0.0]), th_a_out=np.packbits(ref.apply_adaptive_threshold(movie, 51, 0.0)))
    movie = _texture(2, 160, 150, 4)
    payload.update(th_b_movie=movie, th_b_args=np.array([151, -5.0]), th_b_out=np.packbits(ref.apply_adaptive_threshold(movie, 151, -5)))
    movie = _texture(1, 33, 47, 5).astype(np.float32)
    payload.update(th_c_movie=movie, th_c_args=np.array([5, 2.5]), th_c_out=np.packbits(ref.apply_adaptive_threshold(movie, 5, 2.5)))
    # --- CLAHE (tile grid dividing the frame; and the reflected-border path with a clip limit that bites)
    with contextlib.redirect_stdout(io.StringIO()):
        movie = _texture(2, 100, 100, 6) * 150.0
        payload.update(cl_a_movie=movie, cl_a_args=np.array([50000, 10]), cl_a_out=ref.apply_clahe(movie, 50000, 10).astype(np.uint16))
        movie = _texture(2, 90, 70, 7) * 40.0
        payload.update(cl_b_movie=movie, cl_b_args=np.array([1000, 10]), cl_b_out=ref.apply_clahe(movie, 1000, 10).astype(np.uint16))
        movie = np.round(_texture(1, 64, 96, 8))         # few distinct values: clipping and residual redistribution
        payload.update(cl_c_movie=movie, cl_c_args=np.array([2.0, 4]), cl_c_out=ref.apply_clahe(movie, 2.0, 4).astype(np.uint16))
    # --- background subtraction: conduct_optical_flow hands movie_to_analyse back as 'blurred_data' (:195-198, :214)
    movie = _texture(2, 64, 72, 9)
    with contextlib.redirect_stdout(io.StringIO()):
        res = ref.conduct_optical_flow(movie, boxsize=5, background=120.0)
    payload.update(bg_movie=movie, bg_level=np.array(120.0), bg_out=res['blurred_data'])
    np.savez_compressed(os.path.join(GOLDEN, 'legacy_callers.npz'), **payload)
    print('legacy_callers', {k: v.shape for k, v in payload.items() if k.endswith('_out') or k.endswith('_vx')})


def make_large():
    """One frame pair at a BENCH size through the unmodified reference (config d-2: 512^2 actin-like movie, about five
    minutes and 7 GB in SuperLU).  The full fields would be 6 MB; the fixture keeps every 8th pixel, the central 64 x 64
    block and the three outermost rows/columns of each field (the boundary overwrite), which pins a global solve."""
    import time
    movie = synthetic.actin_like_movie(512, 512, 2, seed=1)
    kwargs = dict(speed_alpha=1e5, remodelling_alpha=1e4, smoothing_sigma=2.5)
    t0 = time.perf_counter()
    result, residuals, log = ref_import.reference_solve(movie, **kwargs)
    seconds = time.perf_counter() - t0
    payload = dict(residuals=np.asarray(residuals), reference_seconds=np.array(seconds),
                   movie_checksum=np.array([movie.sum(), (movie ** 2).sum()]),
                   functionals=np.array([result['L1_functional'], result['remodelling_functional'], result['speed_functional']]))
    for key in ('v_x', 'v_y', 'speed', 'remodelling'):
        f = result[key][0]
        payload[key + '_sub8'] = f[::8, ::8].copy()
        payload[key + '_centre'] = f[224:288, 224:288].copy()
        payload[key + '_top'] = f[:3].copy(); payload[key + '_bottom'] = f[-3:].copy()
        payload[key + '_left'] = f[:, :3].copy(); payload[key + '_right'] = f[:, -3:].copy()
        payload[key + '_norm'] = np.array(np.linalg.norm(f))
    for k, v in kwargs.items():
        payload['kw_' + k] = np.array(v, dtype=np.float64)
    np.savez_compressed(os.path.join(GOLDEN, 'actin512_pair0_sample.npz'), **payload)
    phases = [line for line in log.splitlines() if 'took' in line or 'time' in line.lower()]
    print('actin512_pair0_sample: %.1f s in the reference, residual %.2e' % (seconds, max(residuals)))
    print('\n'.join(phases[:12]))


if __name__ == '__main__':
    if '--large-only' in sys.argv:
        make_large()
        sys.exit(0)
    if '--legacy-only' not in sys.argv:
        main()
    make_legacy()
    if '--large' in sys.argv:
        make_large()
