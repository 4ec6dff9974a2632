"""Device-side plumbing: torch tensors as device buffers, every compute step through the C ABI.

PyTorch here only allocates device memory, owns streams and does host<->device copies; all arithmetic
of the hot path happens in ``libofb200.so``.  Shapes: images ``[B, Nx, Ny]``, vectors ``[B, 3, Nx, Ny]``.
"""
import ctypes

import numpy as np
import torch

from . import _lib

_DTYPES = {'float32': (torch.float32, _lib.OFB_F32), 'float64': (torch.float64, _lib.OFB_F64)}
_PRECOND = {'mg': _lib.PRECOND_MG, 'multigrid': _lib.PRECOND_MG, 'bjacobi': _lib.PRECOND_BJACOBI,
            'block_jacobi': _lib.PRECOND_BJACOBI}


def resolve_dtype(dtype):
    if isinstance(dtype, torch.dtype):
        dtype = {torch.float32: 'float32', torch.float64: 'float64'}[dtype]
    dtype = np.dtype(dtype).name if not isinstance(dtype, str) else dtype
    if dtype not in _DTYPES:
        raise ValueError("dtype must be 'float32' or 'float64', got %r" % (dtype,))
    return _DTYPES[dtype]


def _require_cuda(*tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise _lib.OfbError('ofb200 needs CUDA tensors; there is no CPU fallback')


def _stream_ptr(device):
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else None


def gaussian_weights(sigma, truncate=4.0):
    """Normalised Gaussian taps of scipy.ndimage.gaussian_filter1d (radius int(truncate*sigma+0.5))."""
    radius = int(truncate * float(sigma) + 0.5)
    k = np.arange(-radius, radius + 1, dtype=np.float64)
    w = np.exp(-0.5 / (float(sigma) * float(sigma)) * k * k)
    return w / w.sum(), radius


def _code(t):
    return _lib.OFB_F32 if t.dtype == torch.float32 else _lib.OFB_F64


def blur(frames, sigma, out=None):
    """K1: Gaussian blur of ``frames [F,Nx,Ny]`` (mode nearest, truncate 4) -> new tensor."""
    _require_cuda(frames)
    frames = frames.contiguous()
    F, Nx, Ny = frames.shape
    w, radius = gaussian_weights(sigma)
    wd = torch.from_numpy(w).to(frames.device)
    out = torch.empty_like(frames) if out is None else out
    tmp = torch.empty_like(frames)
    _lib.check(_lib.load().ofb_blur(_ptr(frames), _ptr(out), _ptr(tmp), _ptr(wd), radius, F, Nx, Ny, _code(frames),
                                    _stream_ptr(frames.device)))
    return out


def build_rhs(P, C, out=None):
    """K2: right-hand side planes ``[B,3,Nx,Ny]`` from previous / current frames."""
    _require_cuda(P, C)
    B, Nx, Ny = P.shape
    assert P.is_contiguous() and C.is_contiguous() and C.shape == P.shape and C.dtype == P.dtype
    out = torch.empty((B, 3, Nx, Ny), dtype=P.dtype, device=P.device) if out is None else out
    _lib.check(_lib.load().ofb_build_rhs(_ptr(P), _ptr(C), _ptr(out), B, Nx, Ny, _code(P), _stream_ptr(P.device)))
    return out


def _weights(value, B, device):
    if isinstance(value, torch.Tensor):
        t = value.to(device=device, dtype=torch.float64).reshape(-1)
    else:
        t = torch.as_tensor(np.broadcast_to(np.asarray(value, dtype=np.float64), (B,)).copy(), device=device)
    assert t.numel() == B
    return t.contiguous()


def apply_operator(P, x, speed_alpha, remodelling_alpha, eliminated=False):
    """K3: y = A x (reference matrix incl. boundary rows) or y = S x (ghost ring eliminated)."""
    _require_cuda(P, x)
    B, Nx, Ny = P.shape
    assert x.shape == (B, 3, Nx, Ny) and x.is_contiguous() and P.is_contiguous() and x.dtype == P.dtype
    a, b = _weights(speed_alpha, B, P.device), _weights(remodelling_alpha, B, P.device)
    y = torch.empty_like(x)
    fn = _lib.load().ofb_apply_eliminated if eliminated else _lib.load().ofb_apply
    _lib.check(fn(_ptr(P), _ptr(x), _ptr(y), _ptr(a), _ptr(b), B, Nx, Ny, _code(P), _stream_ptr(P.device)))
    return y


def smooth(P, b, x_in, speed_alpha, remodelling_alpha, omega=0.8, rows_per_strip=0, out=None, flags=0):
    """One damped block-Jacobi sweep (the multigrid smoother) -> x_out."""
    _require_cuda(P, b, x_in)
    B, Nx, Ny = P.shape
    a, be = _weights(speed_alpha, B, P.device), _weights(remodelling_alpha, B, P.device)
    out = torch.empty_like(x_in) if out is None else out
    opt = make_options('bjacobi', omega=float(omega), rows_per_strip=int(rows_per_strip), flags=int(flags))
    _lib.check(_lib.load().ofb_smooth(_ptr(P), _ptr(b), _ptr(x_in), _ptr(out), _ptr(a), _ptr(be), ctypes.byref(opt),
                                      B, Nx, Ny, _code(P), _stream_ptr(P.device)))
    return out


def make_options(preconditioner='mg', nu_pre=None, nu_post=None, omega=None, check_every=None, rows_per_strip=None,
                 coarse_sweeps=None, flags=None, omega_pre=None, omega_post=None, max_levels=None):
    opt = _lib.default_options()
    opt.precond = _PRECOND[preconditioner] if isinstance(preconditioner, str) else int(preconditioner)
    for name, value in (('nu_pre', nu_pre), ('nu_post', nu_post), ('omega', omega), ('check_every', check_every),
                        ('rows_per_strip', rows_per_strip), ('coarse_sweeps', coarse_sweeps), ('flags', flags),
                        ('max_levels', max_levels)):
        if value is not None:
            setattr(opt, name, value)
    for name, values in (('omega_pre', omega_pre), ('omega_post', omega_post)):
        if values is None and (omega is not None or nu_pre is not None or nu_post is not None):
            values = ()                                   # an explicit cycle shape / omega replaces the default per-sweep weights
        if values is not None:
            for s in range(4):
                getattr(opt, name)[s] = float(values[s]) if s < len(values) else 0.0
    return opt


class Workspace:
    """Caller-owned solver workspace (one torch byte tensor), reused across calls of the same shape."""

    def __init__(self):
        self.buffer = None

    def get(self, B, Nx, Ny, code, precond, device):
        need = _lib.load().ofb_workspace_bytes(B, Nx, Ny, code, precond)
        if need == 0:
            raise _lib.OfbError(_lib.load().ofb_last_error().decode())
        if self.buffer is None or self.buffer.numel() < need or self.buffer.device != device:
            self.buffer = torch.empty(need, dtype=torch.uint8, device=device)
        return self.buffer, need


_default_workspace = Workspace()


def solve(P, b, x, speed_alpha, remodelling_alpha, rtol=1e-6, max_it=1000, options=None, workspace=None):
    """K4/K5: solve A x = b in place (x: in = initial guess, out = solution).

    Returns ``(iterations int32[B], relres float64[B], converged int32[B])`` as device tensors."""
    _require_cuda(P, b, x)
    B, Nx, Ny = P.shape
    assert b.shape == (B, 3, Nx, Ny) and x.shape == b.shape and b.dtype == P.dtype and x.dtype == P.dtype
    assert P.is_contiguous() and b.is_contiguous() and x.is_contiguous()
    options = options or make_options()
    a, be = _weights(speed_alpha, B, P.device), _weights(remodelling_alpha, B, P.device)
    ws, nbytes = (workspace or _default_workspace).get(B, Nx, Ny, _code(P), options.precond, P.device)
    iters = torch.zeros(B, dtype=torch.int32, device=P.device)
    relres = torch.zeros(B, dtype=torch.float64, device=P.device)
    conv = torch.zeros(B, dtype=torch.int32, device=P.device)
    _lib.check(_lib.load().ofb_solve(_ptr(P), _ptr(b), _ptr(x), _ptr(a), _ptr(be), float(rtol), int(max_it),
                                     ctypes.byref(options), _ptr(ws), nbytes, _ptr(iters), _ptr(relres), _ptr(conv),
                                     B, Nx, Ny, _code(P), _stream_ptr(P.device)))
    return iters, relres, conv


def precondition(P, r, speed_alpha, remodelling_alpha, options=None, workspace=None):
    """One application z = M^-1 r of the solver's preconditioner (diagnostics/tests)."""
    _require_cuda(P, r)
    B, Nx, Ny = P.shape
    options = options or make_options()
    a, be = _weights(speed_alpha, B, P.device), _weights(remodelling_alpha, B, P.device)
    ws, nbytes = (workspace or _default_workspace).get(B, Nx, Ny, _code(P), options.precond, P.device)
    z = torch.zeros_like(r)
    _lib.check(_lib.load().ofb_precondition(_ptr(P), _ptr(r), _ptr(z), _ptr(a), _ptr(be), ctypes.byref(options),
                                            _ptr(ws), nbytes, B, Nx, Ny, _code(P), _stream_ptr(P.device)))
    return z


def epilogue(P, C, x, speed_alpha, remodelling_alpha, scale, out=None):
    """K6: returned fields, functionals and moments.

    -> (vx, vy, speed, gamma [B,Nx,Ny], functionals float64 [B,3], moments float64 [B,4])."""
    _require_cuda(P, C, x)
    B, Nx, Ny = P.shape
    a, be = _weights(speed_alpha, B, P.device), _weights(remodelling_alpha, B, P.device)
    if out is None:
        out = torch.empty((4, B, Nx, Ny), dtype=P.dtype, device=P.device)
    fun = torch.empty((B, 3), dtype=torch.float64, device=P.device)
    mom = torch.empty((B, 4), dtype=torch.float64, device=P.device)
    _lib.check(_lib.load().ofb_epilogue(_ptr(P), _ptr(C), _ptr(x), _ptr(out[0]), _ptr(out[1]), _ptr(out[2]),
                                        _ptr(out[3]), _ptr(fun), _ptr(mom), _ptr(a), _ptr(be), float(scale), B, Nx, Ny, _code(P),
                                        _stream_ptr(P.device)))
    return out[0], out[1], out[2], out[3], fun, mom
