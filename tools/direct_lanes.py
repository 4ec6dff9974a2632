"""Sparse direct solves side by side on one GPU: `lanes` host threads, each with its own workspace and stream, each solving
`pairs` 1024^2 pairs one after the other.  Experiment behind profiles/r02_direct_solve_launches.txt: run against a build with
ND_LANES = 4 sets of handles / side streams per device the solves did not overlap (the device is saturated by the library's panel
kernels), so the library keeps ONE solve per device at a time and with it this script only measures that serialisation.
  python tools/direct_lanes.py [frame] [pairs per lane] [lanes ...]"""
import json, os, sys, threading, time
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
from opticalflow_b200 import engine, synthetic

N = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
per_lane = int(sys.argv[2]) if len(sys.argv) > 2 else 4
lane_counts = [int(v) for v in sys.argv[3:]] or [1, 2, 3, 4]
dev = torch.device('cuda', 0)
movie = synthetic.rotating_shearing_movie(N, N, 5, seed=2)
frames = engine.blur(torch.from_numpy(movie).to(dev).float(), 2.5)
P, C = frames[:-1].contiguous(), frames[1:].contiguous()
b = engine.build_rhs(P, C)
engine._direct_plan(N, N, dev)
ref, _, _ = engine.direct_solve(P[:1], b[:1], 1e3, 1e4, engine.Workspace())
torch.cuda.synchronize()
for lanes in lane_counts:
    spaces = [engine.Workspace() for _ in range(lanes)]
    streams = [torch.cuda.Stream(dev) for _ in range(lanes)]
    out = [None] * lanes

    def work(k, reps):
        torch.cuda.set_device(dev)
        with torch.cuda.stream(streams[k]):
            for r in range(reps):
                q = (k + r) % P.shape[0]
                out[k] = engine.direct_solve(P[q:q + 1], b[q:q + 1], 1e3, 1e4, spaces[k]) + (q,)
            streams[k].synchronize()

    def run(reps):
        th = [threading.Thread(target=work, args=(k, reps)) for k in range(lanes)]
        t0 = time.perf_counter()
        for t in th: t.start()
        for t in th: t.join()
        torch.cuda.synchronize()
        return time.perf_counter() - t0
    run(1)                                    # handles, workspaces
    dt = run(per_lane)
    x0, rr, cv, q = out[0]
    same = bool(q != 0 or torch.equal(x0, ref))
    print(json.dumps(dict(frame=N, lanes=lanes, pairs=lanes * per_lane, seconds=round(dt, 3), seconds_per_pair=round(dt / (lanes * per_lane), 4),
                          relres=[float('%.3g' % o[1].item()) for o in out], converged=[int(o[2].item()) for o in out], lane0_matches_single=same)), flush=True)
    del spaces, out
    torch.cuda.empty_cache()
