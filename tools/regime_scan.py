"""Config d-4's regularisation sweep (speed_alpha = logspace(3, 8, 16), remodelling_alpha 1e4) on 1024^2 d-3 pairs:
iterations and time of the default cycle, the V(2,2)/0.8 fallback and plain block-Jacobi BiCGStab per alpha; mode `dropin`
is what the drop-in does with a batch (multigrid BiCGStab probe, then the sparse direct solve for what it left)."""
import argparse
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from opticalflow_b200 import engine, optical_flow, synthetic  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--n', type=int, default=1024)
    ap.add_argument('--batch', type=int, default=4)
    ap.add_argument('--beta', type=float, default=1e4)
    ap.add_argument('--max-it', type=int, default=300)
    ap.add_argument('--alphas', default='')
    ap.add_argument('--modes', default='default,v22,bjacobi')
    ap.add_argument('--dtype', default='float32')
    a = ap.parse_args()
    dev = torch.device('cuda', 0)
    B, N = a.batch, a.n
    movie = synthetic.rotating_shearing_movie(N, N, B + 1, seed=2).astype(np.float32)
    tdtype, _ = engine.resolve_dtype(a.dtype)
    frames = engine.blur(torch.from_numpy(movie).to(dev).to(tdtype), 2.5)
    P, C = frames[:-1].contiguous(), frames[1:].contiguous()
    b = engine.build_rhs(P, C)
    alphas = [float(v) for v in a.alphas.split(',')] if a.alphas else list(np.logspace(3, 8, 16))
    modes = {'default': engine.make_options('mg'), 'v22': engine.make_options('mg', nu_pre=2, nu_post=2, omega=0.8),
             'bjacobi': engine.make_options('bjacobi')}
    names = a.modes.split(',')
    ws, dws = engine.Workspace(), engine.Workspace()
    for al in alphas:
        line = {'alpha': al, 'beta': a.beta}
        for name in names:
            if name == 'dropin':
                x0 = torch.zeros_like(b)
                for rep in range(2):                      # the first pass allocates the workspaces
                    torch.cuda.synchronize()
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record()
                    x, it, rr, cv, fb, rec = optical_flow._solve_pairs(P, C, al, a.beta, x0, 1e-6, a.max_it, 'mg', ws, {}, False, dws)
                    e1.record()
                    torch.cuda.synchronize()
                line[name] = dict(iters=float(it.float().mean()), conv=int(cv.sum()), direct=int(fb.sum()),
                                  ms_per_pair=round(e0.elapsed_time(e1) / B, 2), relres=float(rr.max()))
                continue
            av = torch.full((B,), al, dtype=torch.float64, device=dev)
            bv = torch.full((B,), a.beta, dtype=torch.float64, device=dev)
            x = torch.zeros_like(b)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            it, rr, cv = engine.solve(P, b, x, av, bv, 1e-6, a.max_it, modes[name])
            e1.record()
            torch.cuda.synchronize()
            line[name] = dict(iters=float(it.float().mean()), conv=int(cv.sum()), ms_per_pair=round(e0.elapsed_time(e1) / B, 2),
                              relres=float(rr.max()))
        print(json.dumps(line), flush=True)


if __name__ == '__main__':
    main()
