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


if __name__ == '__main__':
    main()
