"""BASELINE.json's configs d-1, d-2 and d-4 end to end through the drop-in (host numpy in, host numpy out), one JSON line
each.  d-3 (the headline) and d-5 are bench.py's workloads.

  python tools/config_bench.py [--d4-pairs 8] [--skip d-4]

d-1: the 256^2 translated pair, checked against the oracle's direct solve (the reference's own CPU-runnable case).
d-2: the 100-frame 512^2 actin-like movie (99 pairs), float32, batched on one GPU.
d-4: 16 speed_alpha values logspace(3, 8, 16) x `--d4-pairs` pairs of the 1024^2 rotating+shearing movie through
     vary_regularisation (the full config has 200 pairs; cells scale linearly in the pair count)."""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from opticalflow_b200 import optical_flow, synthetic  # noqa: E402


def timed(fn, reps=2):
    fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        out = fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps, out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--d4-pairs', type=int, default=8)
    ap.add_argument('--skip', default='')
    ap.add_argument('--d4-fast-only', action='store_true', help='d-4: only the cells the iterative solver converges')
    a = ap.parse_args()
    kw = dict(speed_alpha=1e5, remodelling_alpha=1e4, smoothing_sigma=2.5)
    if 'd-1' not in a.skip:
        from oracle import flow_oracle
        movie = synthetic.translating_movie(256, 256, 2)
        s, res = timed(lambda: optical_flow.variational_optical_flow(movie, dtype='float64', rtol=1e-10, **kw))
        t0 = time.perf_counter()
        want = flow_oracle.variational_optical_flow(movie, **kw)
        cpu_s = time.perf_counter() - t0
        err = {k: float(np.linalg.norm(res[k] - want[k]) / np.linalg.norm(want[k])) for k in ('v_x', 'v_y', 'remodelling')}
        inner = (slice(None), slice(8, -8), slice(8, -8))
        print(json.dumps(dict(config='d-1', shape=[2, 256, 256], dtype='float64', rtol=1e-10, seconds=round(s, 4),
                              pairs_per_s=round(1 / s, 1), iterations=res['solver_info']['iterations'].tolist(),
                              rel_l2_vs_oracle_direct_solve=err, oracle_cpu_seconds=round(cpu_s, 1),
                              mean_v=[float(res['v_x'][inner].mean()), float(res['v_y'][inner].mean())])), flush=True)
    if 'd-2' not in a.skip:
        movie = synthetic.actin_like_movie(512, 512, 100, seed=1).astype(np.float32)
        s, res = timed(lambda: optical_flow.variational_optical_flow(movie, dtype='float32', **kw))
        it = res['solver_info']['iterations']
        print(json.dumps(dict(config='d-2', shape=list(movie.shape), dtype='float32', rtol=1e-6, seconds=round(s, 4),
                              pairs_per_s=round(99 / s, 1), iterations_mean=float(it.mean()), iterations_max=int(it.max()),
                              converged=int(res['solver_info']['converged'].sum()), fallback=int(res['solver_info']['fallback'].sum()),
                              mean_v=[float(res['v_x'].mean()), float(res['v_y'].mean())], imposed_v=[0.8, 0.5])), flush=True)
    if 'd-4' not in a.skip:
        n = a.d4_pairs
        import bench                                  # the pool-of-processes generator of the d-3 movie
        movie = bench.generate_movie(n + 1, 2)
        alphas = np.logspace(3, 8, 16)
        if a.d4_fast_only:
            alphas = alphas[alphas >= 4e4]
        optical_flow.vary_regularisation(movie[:3], alphas, np.array([1e4]), dtype='float32', smoothing_sigma=2.5, max_it=300)   # warm-up
        t0 = time.perf_counter()
        res = optical_flow.vary_regularisation(movie, alphas, np.array([1e4]), dtype='float32', smoothing_sigma=2.5, max_it=300)
        torch.cuda.synchronize()
        s = time.perf_counter() - t0
        conv = res['converged'][:, 0]
        print(json.dumps(dict(config='d-4 (%d of 200 pairs)' % n, cells=len(alphas), pairs=n, dtype='float32', rtol=1e-6, max_it=300,
                              seconds=round(s, 3), solves_per_s=round(len(alphas) * n / s, 1), converged_cells=int(conv.sum()),
                              converged=conv.astype(int).tolist(), speed_means=[float('%.4g' % v) for v in res['speed_means'][:, 0]])),
              flush=True)
        if a.d4_fast_only:
            return
        hi = alphas >= 4e4                      # the cells the multigrid cycle converges (DESIGN.md regime map)
        s, res = timed(lambda: optical_flow.vary_regularisation(movie, alphas[hi], np.array([1e4]), dtype='float32', smoothing_sigma=2.5))
        print(json.dumps(dict(config='d-4, the %d cells with alpha >= 4.6e4' % int(hi.sum()), pairs=n, seconds=round(s, 3),
                              solves_per_s=round(int(hi.sum()) * n / s, 1), converged_cells=int(res['converged'].sum()))), flush=True)


if __name__ == '__main__':
    main()
