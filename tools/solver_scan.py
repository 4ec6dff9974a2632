"""Scan multigrid cycle parameters on the bench workload (d-3, 1024^2 fp32): iterations and ms per batch."""
import argparse
import itertools
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from opticalflow_b200 import engine, synthetic  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--n', type=int, default=1024)
    ap.add_argument('--batch', type=int, default=8)
    ap.add_argument('--dtype', default='float32')
    ap.add_argument('--rtol', type=float, default=1e-6)
    ap.add_argument('--pre', default='1,2')
    ap.add_argument('--post', default='1,2')
    ap.add_argument('--omega', default='0.7,0.8,0.9')
    ap.add_argument('--workload', default='rot')
    ap.add_argument('--cycles', default='', help='explicit cycles: "wpre0,wpre1|wpost0,wpost1;..." (per-sweep damping)')
    ap.add_argument('--max-levels', type=int, default=0)
    ap.add_argument('--coarse-sweeps', type=int, default=30)
    ap.add_argument('--max-it', type=int, default=200)
    ap.add_argument('--alpha', type=float, default=1e5)
    ap.add_argument('--beta', type=float, default=1e4)
    a = ap.parse_args()
    dev = torch.device('cuda', 0)
    tdtype, _ = engine.resolve_dtype(a.dtype)
    B, N = a.batch, a.n
    if a.workload == 'rot':
        movie = synthetic.rotating_shearing_movie(N, N, B + 1, seed=2)
    elif a.workload == 'actin':
        movie = synthetic.actin_like_movie(N, N, B + 1)
    else:
        movie = synthetic.translating_movie(N, N, B + 1)
    frames = engine.blur(torch.from_numpy(movie).to(dev).to(tdtype), 2.5)
    P, C = frames[:-1], frames[1:]
    b = engine.build_rhs(P, C)
    alpha = torch.full((B,), a.alpha, dtype=torch.float64, device=dev)
    beta = torch.full((B,), a.beta, dtype=torch.float64, device=dev)
    ws = engine.Workspace()
    configs = []
    if a.cycles:
        for c in a.cycles.split(';'):
            pre_w, post_w = c.split('|')
            pre_w = [float(v) for v in pre_w.split(',') if v]
            post_w = [float(v) for v in post_w.split(',') if v]
            configs.append((len(pre_w), len(post_w), pre_w, post_w))
    else:
        for pre, post, om in itertools.product([int(v) for v in a.pre.split(',')], [int(v) for v in a.post.split(',')],
                                               [float(v) for v in a.omega.split(',')]):
            configs.append((pre, post, [om] * pre, [om] * post))
    for pre, post, pre_w, post_w in configs:
        om = pre_w[0]
        opts = engine.make_options('mg', nu_pre=pre, nu_post=post, omega=om, omega_pre=pre_w, omega_post=post_w,
                                   max_levels=a.max_levels, coarse_sweeps=a.coarse_sweeps)
        times = []
        for rep in range(3):
            x = torch.zeros_like(b)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            it, rr, cv = engine.solve(P, b, x, alpha, beta, a.rtol, a.max_it, opts, ws)
            e1.record()
            torch.cuda.synchronize()
            times.append(e0.elapsed_time(e1))
        print(json.dumps(dict(pre=pre_w, post=post_w, iters=float(it.float().mean()), iters_max=int(it.max()),
                              converged=int(cv.sum()), ms=round(min(times), 2), ms_per_pair=round(min(times) / B, 3))), flush=True)


if __name__ == '__main__':
    main()
