"""Sparse direct solve (ofb_direct_solve) on d-3 frame pairs: time, workspace and true residual per frame size and alpha.
  python tools/direct_bench.py [sizes...]"""
import json, os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
from opticalflow_b200 import engine, synthetic

sizes = [int(v) for v in sys.argv[1:]] or [256, 512, 1024]
dev = torch.device('cuda', 0)
for N in sizes:
    movie = synthetic.rotating_shearing_movie(N, N, 3, seed=2)
    for dtype in (torch.float64, torch.float32):
        frames = engine.blur(torch.from_numpy(movie).to(dev).to(dtype), 2.5)
        P, C = frames[:-1].contiguous(), frames[1:].contiguous()
        b = engine.build_rhs(P, C)
        ws = engine.Workspace()
        t0 = time.perf_counter(); engine._direct_plan(N, N, dev); plan_s = time.perf_counter() - t0
        for alpha in (1e5, 1e3):
            x, rr, cv = engine.direct_solve(P, b, alpha, 1e4, ws)        # warm-up (handles, workspace)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            x, rr, cv = engine.direct_solve(P, b, alpha, 1e4, ws)
            torch.cuda.synchronize()
            dt = (time.perf_counter() - t0) / P.shape[0]
            line = dict(frame=N, dtype=str(dtype).replace('torch.', ''), alpha=alpha, seconds_per_pair=round(dt, 4), relres=[float('%.3g' % v) for v in rr.tolist()],
                        workspace_gb=round(ws.buffer.numel() / 1e9, 2), plan_seconds=round(plan_s, 2),
                        graph=int(__import__('opticalflow_b200._lib', fromlist=['x']).load().ofb_direct_graph_state()))
            if alpha == 1e5:                                               # where the Krylov path works: the two must agree
                xk = torch.zeros_like(b)
                it, rk, cv = engine.solve(P, b, xk, alpha, 1e4, rtol=1e-10 if dtype == torch.float64 else 1e-6, max_it=200)
                line['rel_l2_vs_multigrid_bicgstab'] = float(((x - xk).norm() / xk.norm()).item())
                line['krylov_iterations'] = it.tolist()
            print(json.dumps(line), flush=True)
        del ws, x, b, P, C, frames
        torch.cuda.empty_cache()
