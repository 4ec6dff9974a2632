"""Micro-benchmark of single kernels through the C ABI (CUDA events, inputs larger than L2).

  python tools/kernel_bench.py [--n 1024] [--batch 16] [--dtype float32]

Prints one line per variant: kernel, ms per launch, algorithmic GB/s and fraction of MEASURED_PEAKS.json hbm_gbs."""
import argparse
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from opticalflow_b200 import _lib, engine  # noqa: E402


def timeit(fn, reps=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--n', type=int, default=1024)
    ap.add_argument('--nx', type=int, default=0, help='rows (default --n)')
    ap.add_argument('--batch', type=int, default=16)
    ap.add_argument('--dtype', default='float32')
    ap.add_argument('--rows', default='0,8,16,24,32,48,64,128')
    a = ap.parse_args()
    dev = torch.device('cuda', 0)
    tdtype, _ = engine.resolve_dtype(a.dtype)
    w = 4 if a.dtype == 'float32' else 8
    B, N = a.batch, a.n
    NX = a.nx or N
    peak = 6650.0
    pf = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.isfile(pf):
        peak = float(json.load(open(pf))['hbm_gbs'])
    g = torch.Generator(device=dev).manual_seed(0)
    P = (torch.rand((B, NX, N), device=dev, generator=g, dtype=tdtype) * 100 + 50).contiguous()
    b = torch.randn((B, 3, NX, N), device=dev, generator=g, dtype=tdtype)
    x = torch.randn((B, 3, NX, N), device=dev, generator=g, dtype=tdtype)
    y = torch.empty_like(x)
    alpha = torch.full((B,), 1e5, dtype=torch.float64, device=dev)
    beta = torch.full((B,), 1e4, dtype=torch.float64, device=dev)
    px = B * NX * N

    def report(name, ms, words):
        gbs = words * w * px / (ms * 1e-3) / 1e9
        print(json.dumps(dict(kernel=name, ms=round(ms, 4), alg_gbs=round(gbs, 1), frac=round(gbs / peak, 3))), flush=True)

    ms = timeit(lambda: y.copy_(x))
    report('torch copy (3 planes r+w)', ms, 6)
    for flags, tag in ((0, 'stream'), (_lib.FLAG_TILE_FASTEST, 'stream tile-fastest'), (_lib.FLAG_NO_STREAM, 'register')):
        for rows in [int(r) for r in a.rows.split(',')]:
            ms = timeit(lambda: engine.smooth(P, b, x, alpha, beta, rows_per_strip=rows, out=y, flags=flags))
            report('smooth %s rows=%d' % (tag, rows), ms, 10)
    for flags, tag in ((0, 'stream'),):
        ms = timeit(lambda: engine.apply_operator(P, x, alpha, beta, eliminated=True))
        report('apply_eliminated %s (+alloc)' % tag, ms, 7)


if __name__ == '__main__':
    main()
