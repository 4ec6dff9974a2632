"""A/B of the fused prolongation (OFB_FLAG_FUSED_PROLONG): one preconditioner application on 64 pairs of 1024^2 fp32."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from opticalflow_b200 import _lib, engine  # noqa: E402

dev = torch.device('cuda', 0)
B = 64
g = torch.Generator(device=dev).manual_seed(0)
P = (torch.rand((B, 1024, 1024), device=dev, generator=g) * 150 + 30).contiguous()
r = torch.randn((B, 3, 1024, 1024), device=dev, generator=g)
r[:, :, 0, :] = 0; r[:, :, -1, :] = 0; r[:, :, :, 0] = 0; r[:, :, :, -1] = 0
if os.environ.get('OFB_LIB'):
    _lib.LIBRARY_PATH = os.environ['OFB_LIB']
lib = _lib.load()
for name, flags in (('fused', _lib.FLAG_FUSED_PROLONG), ('plain', 0), ('fused', _lib.FLAG_FUSED_PROLONG), ('plain', 0)):
    opt = engine.make_options('mg', flags=flags)
    engine.precondition(P, r, 1e5, 1e4, opt)
    torch.cuda.synchronize()
    n0 = lib.ofb_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        z = engine.precondition(P, r, 1e5, 1e4, opt)
    e1.record(); torch.cuda.synchronize()
    print(name, 'launches per application', (lib.ofb_launch_count() - n0) / 5, 'ms', e0.elapsed_time(e1) / 5, float(z.double().norm()), flush=True)
