"""One warm sparse direct solve of a d-3 pair for the profiler:  ncu ... python tools/direct_profile.py [N]"""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
from opticalflow_b200 import engine, synthetic
N = int(sys.argv[1]) if len(sys.argv) > 1 else 512
dev = torch.device('cuda', 0)
frames = engine.blur(torch.from_numpy(synthetic.rotating_shearing_movie(N, N, 2, seed=2)).to(dev), 2.5)
P, C = frames[:1].contiguous(), frames[1:].contiguous()
b = engine.build_rhs(P, C)
ws = engine.Workspace()
engine.direct_solve(P, b, 1e3, 1e4, ws)        # warm-up: handles, workspace, plan
torch.cuda.synchronize()
torch.cuda.profiler.start()
x, rr, cv = engine.direct_solve(P, b, 1e3, 1e4, ws)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print('relres', rr.tolist(), 'converged', cv.tolist())
