"""Host timeline of the drop-in on the bench workload (129-frame 1024^2 movie, fp32):  python tools/e2e_profile.py [B]"""
import os, sys, time, json
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
from opticalflow_b200 import optical_flow, synthetic
B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
movie = synthetic.rotating_shearing_movie(1024, 1024, 17, seed=2).astype(np.float32)
zig = list(range(17))
while len(zig) < 129:
    zig += list(range(15, -1, -1)) + list(range(1, 17))
m = np.ascontiguousarray(movie[np.array(zig[:129])])
kw = dict(speed_alpha=1e5, remodelling_alpha=1e4, smoothing_sigma=2.5, dtype='float32', rtol=1e-6, batch_pairs=B)
res = None
for rep in range(6):
    t = time.perf_counter(); res = optical_flow.variational_optical_flow(m, **kw); dt = time.perf_counter() - t
    tl = res['solver_info']['timeline']
    print('rep %d call %.1f ms -> %.1f pairs/s | chunks (wait,solve): %s | %s | egress %s' % (
        rep, dt * 1e3, (m.shape[0] - 1) / dt, [(round(r['wait_slot_and_frames'] * 1e3, 1), round(r['solve'] * 1e3, 1)) for r in tl[:-1]],
        {k: round(v * 1e3, 1) for k, v in tl[-1].items()},
        [(round(a * 1e3, 1), round(b * 1e3, 1)) for a, b in res['solver_info']['egress_log']]), flush=True)
    if rep == 3:
        del res; res = None; import gc; gc.collect()
