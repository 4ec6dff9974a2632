"""Micro-benchmark of the NEXT-3 / NEXT-4 kernels through the C ABI (CUDA events, inputs larger than L2), with the numpy
restatement timed beside them on a bounded sample.

  python tools/legacy_bench.py [--n 1024] [--batch 16] [--cpu]

One JSON line per kernel: ms per launch, algorithmic GB/s and fraction of MEASURED_PEAKS.json hbm_gbs.
Algorithmic bytes (w = bytes per scalar): Liu-Shen sweep 6w per pixel (read vx, vy, P, C; write vx', vy');
adaptive threshold w + 1 (read movie, write bool; the uint8 plane and the window re-reads are L2 traffic);
CLAHE 2 passes over the frame + one result plane = 3w (histograms and look-up tables live in L2)."""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from opticalflow_b200 import engine  # noqa: E402
from tools.kernel_bench import timeit  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--n', type=int, default=1024)
    ap.add_argument('--batch', type=int, default=16)
    ap.add_argument('--cpu', action='store_true', help='also time the numpy restatement on one frame / pair')
    a = ap.parse_args()
    dev = torch.device('cuda', 0)
    B, N = a.batch, a.n
    peak = 6650.0
    pf = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.isfile(pf):
        peak = float(json.load(open(pf))['hbm_gbs'])
    g = torch.Generator(device=dev).manual_seed(0)
    px = B * N * N

    def report(name, ms, bytes_per_px, **extra):
        gbs = bytes_per_px * px / (ms * 1e-3) / 1e9
        print(json.dumps(dict(kernel=name, ms=round(ms, 4), alg_gbs=round(gbs, 1), frac=round(gbs / peak, 3), frames=B, n=N, **extra)),
              flush=True)

    for name, tdtype, w in (('float32', torch.float32, 4), ('float64', torch.float64, 8)):
        frames = (torch.rand((B + 1, N, N), device=dev, generator=g, dtype=tdtype) * 200 + 20).contiguous()
        vx = torch.randn((B, N, N), device=dev, generator=g, dtype=tdtype) * 0.1
        vy = torch.randn((B, N, N), device=dev, generator=g, dtype=tdtype) * 0.1
        its = 10
        ms = timeit(lambda: engine.liu_shen(frames[:B], frames[1:], vx, vy, 2e3, its), reps=5, warm=2)
        # 10 sweeps + the finish pass (read 2, write 3 planes): report per sweep
        report('liu_shen sweep %s' % name, ms / (its + 5.0 / 6.0), 6 * w, launches_per_call=its + 1)
        ms = timeit(lambda: engine.adaptive_threshold(frames[:B], 51, 0.0), reps=5, warm=2)
        report('adaptive_threshold k=51 %s' % name, ms, w + 1)
        ms = timeit(lambda: engine.adaptive_threshold(frames[:B], 151, -5), reps=5, warm=2)
        report('adaptive_threshold k=151 %s' % name, ms, w + 1)
        movie16 = frames[:B] * 200.0
        ms = timeit(lambda: engine.clahe(movie16, 50000, 10, 10, frames_in_flight=B), reps=5, warm=2)
        report('clahe 10x10 tiles %s' % name, ms, 3 * w)
        blurred = engine.blur(frames[:B], 10)
        ms = timeit(lambda: engine.subtract_background(frames[:B], blurred, 120.0), reps=10, warm=2)
        report('subtract_background %s' % name, ms, 3 * w)
        del frames, vx, vy, movie16, blurred
    if a.cpu:
        from oracle import legacy_oracle as lo
        rng = np.random.default_rng(0)
        movie = rng.uniform(20, 220, (2, N, N))
        t0 = time.perf_counter()
        lo.liu_shen_optical_flow(movie, 1.0, 1.0, 2e3, 1.0, None, None, None, 10)
        print(json.dumps(dict(cpu='numpy liu_shen 10 sweeps, 1 pair', n=N, ms_per_sweep=round((time.perf_counter() - t0) * 100, 2))))
        t0 = time.perf_counter()
        lo.apply_adaptive_threshold(movie[:1], 51, 0.0)
        print(json.dumps(dict(cpu='numpy adaptive_threshold k=51, 1 frame', n=N, ms=round((time.perf_counter() - t0) * 1e3, 2))))
        t0 = time.perf_counter()
        lo.apply_clahe(movie[:1] * 200, 50000, 10)
        print(json.dumps(dict(cpu='numpy clahe, 1 frame', n=N, ms=round((time.perf_counter() - t0) * 1e3, 2))))


if __name__ == '__main__':
    main()
