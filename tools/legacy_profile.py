"""Runs a few Liu-Shen sweeps (and optionally the other NEXT-3 kernels) for an ncu capture:
  ncu --set full --clock-control none --import-source on -k regex:k_liu_shen_step -s 4 -c 1 -o out python tools/legacy_profile.py"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from opticalflow_b200 import engine  # noqa: E402

dtype = torch.float64 if '--f64' in sys.argv else torch.float32
dev = torch.device('cuda', 0)
B, N = 16, 1024
g = torch.Generator(device=dev).manual_seed(0)
frames = (torch.rand((B + 1, N, N), device=dev, generator=g, dtype=dtype) * 200 + 20).contiguous()
vx = torch.randn((B, N, N), device=dev, generator=g, dtype=dtype) * 0.1
vy = torch.randn((B, N, N), device=dev, generator=g, dtype=dtype) * 0.1
engine.liu_shen(frames[:B], frames[1:], vx, vy, 2e3, 8)
if '--all' in sys.argv:
    engine.adaptive_threshold(frames[:B], 151, -5)
    engine.clahe(frames[:B] * 200.0, 50000, 10, 10, frames_in_flight=B)
torch.cuda.synchronize()
