"""warm_start=True (the reference's seeding of a pair with an earlier solution, optical_flow.py:803-806, here across chunks):
iterations per pair with and without, on the stationary flow of config d-3 and on a flow whose rotation rate changes in time.
  python tools/warm_start_bench.py [frames] [N]"""
import json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
from opticalflow_b200 import optical_flow, synthetic

T = int(sys.argv[1]) if len(sys.argv) > 1 else 65
N = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
base = synthetic.textured_frame(N, N, seed=2)
ii, jj = np.meshgrid(np.arange(N, dtype=np.float64), np.arange(N, dtype=np.float64), indexing='ij')
c = (N - 1) / 2.0


def movie(omega_of_t):
    """frame t = frame 0 rotated by the accumulated angle (plus the d-3 shear): the pair (t, t+1) sees the rate omega(t)."""
    frames, angle = [base], 0.0
    for t in range(1, T):
        angle += omega_of_t(t - 1)
        vx = -angle * (jj - c) + 5e-4 * t * (jj - c)
        vy = angle * (ii - c)
        frames.append(synthetic._bicubic_sample(base, ii - vx, jj - vy))
    return np.stack(frames).astype(np.float32)


kw = dict(speed_alpha=1e5, remodelling_alpha=1e4, smoothing_sigma=2.5, dtype='float32', rtol=1e-6, batch_pairs=16)
for name, om in (('stationary (config d-3: omega 1e-3 per frame)', lambda t: 1e-3),
                 ('rotation rate swinging between -1e-3 and 1e-3 with period 24 frames', lambda t: 1e-3 * np.cos(2 * np.pi * t / 24.0))):
    m = movie(om)
    cold = optical_flow.variational_optical_flow(m, **kw)
    warm = optical_flow.variational_optical_flow(m, warm_start=True, **kw)
    err = float(np.linalg.norm(warm['v_x'] - cold['v_x']) / np.linalg.norm(cold['v_x']))
    print(json.dumps(dict(flow=name, frame=N, pairs=T - 1, chunk=16, iterations_cold=float(cold['solver_info']['iterations'].mean()),
                          iterations_warm=float(warm['solver_info']['iterations'].mean()),
                          iterations_warm_after_first_chunk=float(warm['solver_info']['iterations'][16:].mean()),
                          rel_l2_warm_vs_cold=err, seconds_cold=cold['solver_info']['elapsed_seconds'],
                          seconds_warm=warm['solver_info']['elapsed_seconds'])), flush=True)
