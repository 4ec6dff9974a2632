"""Host timeline of the drop-in on a long d-3 movie (1024^2, fp32):  python tools/e2e_profile.py [frames] [lean]

Frames are cycled from 65 generated ones (zig-zag: every pair is a valid d-3 pair); prints, per call, the per-chunk waits
and solve times, the set-up / statistics / drain times and the egress worker's log."""
import gc, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
from opticalflow_b200 import optical_flow, synthetic
T = int(sys.argv[1]) if len(sys.argv) > 1 else 500
lean = len(sys.argv) > 2
movie = synthetic.rotating_shearing_movie(1024, 1024, 17, seed=2).astype(np.float32)
zig = list(range(17))
while len(zig) < T:
    zig += list(range(15, -1, -1)) + list(range(1, 17))
m = np.ascontiguousarray(movie[np.array(zig[:T])])
kw = dict(speed_alpha=1e5, remodelling_alpha=1e4, smoothing_sigma=2.5, dtype='float32', rtol=1e-6)
if lean:
    kw.update(result_dtype='float32', outputs=('v_x', 'v_y', 'remodelling'))
for rep in range(4):
    t = time.perf_counter(); res = optical_flow.variational_optical_flow(m, **kw); dt = time.perf_counter() - t
    info = res['solver_info']; tl = info['timeline']
    del res; gc.collect()
    ms = lambda v: round(v * 1e3, 1)
    print('rep %d call %.1f ms -> %.1f pairs/s | chunks (wait, prepare, solve, submit): %s | %s | egress (wait, work) %s' % (
        rep, dt * 1e3, (T - 1) / dt, [(ms(r['wait_slot_and_frames']), ms(r['prepare']), ms(r['solve']), ms(r['submit'])) for r in tl[:-1]],
        {k: ms(v) for k, v in tl[-1].items()}, [(ms(a), ms(b)) for a, b in info['egress_log']]), flush=True)
