"""GPU test of the allocation-free entry point ``ofb_blur_noalloc`` (include/ofb200.h; no counterpart in the
reference).  Collected last on purpose
(``zz``).

The new entry point is called straight through the C ABI (ctypes) and held to the existing ``ofb_blur`` on the same
inputs - within the tolerances tests/test_gpu_parity.py uses for the blur (1e-10 absolute in float64, 2e-3 in float32
on 0..255 data) - on two calls in a row with different shapes and kernels, so that what the first call left in the
shared scratch array cannot leak into the second."""
import ctypes

import numpy as np
import pytest
import torch

from opticalflow_b200 import _lib, engine

pytestmark = pytest.mark.gpu

TOL = {torch.float64: 1e-10, torch.float32: 2e-3}


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr())


def _stream(device):
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def _blur_both(lib, frames, sigma):
    """(existing ofb_blur, new ofb_blur_noalloc) on the same device frames; every buffer is the caller's."""
    F, Nx, Ny = frames.shape
    w, radius = engine.gaussian_weights(sigma)
    wd = torch.from_numpy(w).to(frames.device)
    code = _lib.OFB_F32 if frames.dtype == torch.float32 else _lib.OFB_F64
    want, tmp, got = torch.empty_like(frames), torch.empty_like(frames), torch.full_like(frames, float('nan'))
    _lib.check(lib.ofb_blur(_ptr(frames), _ptr(want), _ptr(tmp), _ptr(wd), radius, F, Nx, Ny, code, _stream(frames.device)))
    torch.cuda.synchronize(frames.device)
    free_before = torch.cuda.mem_get_info(frames.device)[0]
    _lib.check(lib.ofb_blur_noalloc(_ptr(frames), _ptr(got), _ptr(wd), radius, F, Nx, Ny, code, _stream(frames.device)))
    torch.cuda.synchronize(frames.device)
    free_after = torch.cuda.mem_get_info(frames.device)[0]
    return want, got, free_before - free_after


@pytest.mark.parametrize('dtype', [torch.float64, torch.float32])
def test_blur_noalloc_matches_blur_on_two_calls_in_a_row(cuda_device, dtype):
    lib = _lib.load()
    rng = np.random.default_rng(7)
    with torch.cuda.device(cuda_device):
        # first call: ragged shape, wide kernel; second call: another shape and kernel over the same scratch
        cases = [((3, 150, 300), 5.0), ((2, 257, 132), 1.3), ((3, 150, 300), 5.0)]
        for n, (shape, sigma) in enumerate(cases):
            frames = torch.from_numpy(rng.uniform(0, 255, shape)).to(cuda_device).to(dtype).contiguous()
            want, got, grown = _blur_both(lib, frames, sigma)
            assert torch.isfinite(got).all()
            assert float((got - want).abs().max()) < TOL[dtype], (n, shape, sigma)
            if n > 0:                         # the first call may load the module lazily; after that nothing is allocated
                assert grown == 0, grown


def test_blur_noalloc_rejects_what_does_not_fit(cuda_device):
    lib = _lib.load()
    with torch.cuda.device(cuda_device):
        w, radius = engine.gaussian_weights(2.0)
        wd = torch.from_numpy(w).to(cuda_device)
        # exactly the limit passes: 4 float32 frames of 1024^2 = 16 MiB
        frames = torch.rand((4, 1024, 1024), dtype=torch.float32, device=cuda_device)
        out = torch.empty_like(frames)
        assert frames.numel() * 4 == _lib.BLUR_NOALLOC_MAX_BYTES
        _lib.check(lib.ofb_blur_noalloc(_ptr(frames), _ptr(out), _ptr(wd), radius, 4, 1024, 1024, _lib.OFB_F32, _stream(cuda_device)))
        want = engine.blur(frames, 2.0)
        torch.cuda.synchronize(cuda_device)
        assert float((out - want).abs().max()) < 2e-3 / 255          # data in 0..1 here
        # one row more does not
        big = torch.rand((4, 1025, 1024), dtype=torch.float32, device=cuda_device)
        out = torch.empty_like(big)
        rc = lib.ofb_blur_noalloc(_ptr(big), _ptr(out), _ptr(wd), radius, 4, 1025, 1024, _lib.OFB_F32, _stream(cuda_device))
        assert rc == -12 and b'OFB_BLUR_NOALLOC_MAX_BYTES' in lib.ofb_last_error()
        # the same frames in float64 need twice the bytes
        rc = lib.ofb_blur_noalloc(_ptr(frames), _ptr(out), _ptr(wd), radius, 4, 1024, 1024, _lib.OFB_F64, _stream(cuda_device))
        assert rc == -12
        rc = lib.ofb_blur_noalloc(_ptr(frames), _ptr(frames), _ptr(wd), radius, 4, 1024, 1024, _lib.OFB_F32, _stream(cuda_device))
        assert rc < 0 and b'alias' in lib.ofb_last_error()
