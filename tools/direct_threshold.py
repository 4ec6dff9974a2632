"""One 1024^2 float pair, alpha 1e3, by the sparse direct solve: seconds per pair (OFB_ND_BATCHED_MAX is read by the library once per process)."""
import json, os, sys, time
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
from opticalflow_b200 import engine, synthetic
N = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
dev = torch.device('cuda', 0)
movie = synthetic.rotating_shearing_movie(N, N, 3, seed=2)
frames = engine.blur(torch.from_numpy(movie).to(dev).float(), 2.5)
P, C = frames[:-1].contiguous(), frames[1:].contiguous()
b = engine.build_rhs(P, C)
ws = engine.Workspace()
engine.direct_solve(P, b, 1e3, 1e4, ws); torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(3): x, rr, cv = engine.direct_solve(P, b, 1e3, 1e4, ws)
torch.cuda.synchronize()
print(json.dumps(dict(frame=N, batched_max=os.environ.get('OFB_ND_BATCHED_MAX', 'default'), seconds_per_pair=round((time.perf_counter() - t0) / (3 * P.shape[0]), 4),
                      relres=[float('%.3g' % v) for v in rr.tolist()], converged=cv.tolist())))
