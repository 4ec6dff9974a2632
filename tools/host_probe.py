"""Host-side probe for the egress pipeline: THP mode, D2H bandwidth, float32->float64 widening speed into fresh
and into pre-touched numpy arrays, page-fault cost."""
import os
import time

import numpy as np
import torch

print('cpus', os.cpu_count(), 'torch threads', torch.get_num_threads())
for f in ('enabled', 'defrag', 'shmem_enabled'):
    try:
        print('THP', f, open('/sys/kernel/mm/transparent_hugepage/' + f).read().strip())
    except Exception as e:
        print('THP', f, 'n/a', e)
n = 16 * 4 * 1024 * 1024          # one chunk: 16 pairs x 4 planes x 1 Mpx
dev = torch.device('cuda', 0)
src_dev = torch.randn(n, device=dev)
pin = torch.empty(n, dtype=torch.float32, pin_memory=True)
for _ in range(2):
    pin.copy_(src_dev, non_blocking=True); torch.cuda.synchronize()
t = time.perf_counter(); pin.copy_(src_dev, non_blocking=True); torch.cuda.synchronize(); dt = time.perf_counter() - t
print('D2H pinned %.1f MB: %.1f ms, %.1f GB/s' % (n * 4 / 1e6, dt * 1e3, n * 4 / dt / 1e9))
for label, make in (('fresh np.empty', lambda: np.empty(n, np.float64)), ('fresh np.zeros', lambda: np.zeros(n, np.float64))):
    ts = []
    for _ in range(3):
        dst = make()
        t = time.perf_counter(); torch.from_numpy(dst).copy_(pin); ts.append(time.perf_counter() - t)
    print('widen f32->f64 into %s: %.1f ms (%.1f GB/s written)' % (label, min(ts) * 1e3, n * 8 / min(ts) / 1e9))
dst = np.empty(n, np.float64); dst[:] = 0
ts = []
for _ in range(3):
    t = time.perf_counter(); torch.from_numpy(dst).copy_(pin); ts.append(time.perf_counter() - t)
print('widen f32->f64 into pre-touched array: %.1f ms (%.1f GB/s written)' % (min(ts) * 1e3, n * 8 / min(ts) / 1e9))
for th in (1, 2, 4, 8, 16):
    torch.set_num_threads(th)
    ts = []
    for _ in range(3):
        t = time.perf_counter(); torch.from_numpy(dst).copy_(pin); ts.append(time.perf_counter() - t)
    print('  threads %2d: pre-touched %.1f ms' % (th, min(ts) * 1e3))
t = time.perf_counter(); a = np.empty(n, np.float64); a[::512] = 1.0; dt = time.perf_counter() - t
print('first touch of %.0f MB (one write per page): %.1f ms' % (n * 8 / 1e6, dt * 1e3))
t = time.perf_counter(); b32 = np.empty(n, np.float32); torch.from_numpy(b32).copy_(pin); dt = time.perf_counter() - t
print('copy f32->f32 into fresh array: %.1f ms' % (dt * 1e3))
