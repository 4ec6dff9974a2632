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


_weight_cache = {}


def _device_weights(sigma, device):
    """Gaussian taps on the device, uploaded once per (sigma, device): a pageable host->device copy per call would wait
    on the copy engine behind the frame prefetch of the next chunk."""
    key = (float(sigma), str(device))
    hit = _weight_cache.get(key)
    if hit is None:
        w, radius = gaussian_weights(sigma)
        if len(_weight_cache) > 64:
            _weight_cache.clear()
        hit = _weight_cache[key] = (torch.from_numpy(w).to(device), radius)
    return hit


def _code(t):
    return _lib.OFB_F32 if t.dtype == torch.float32 else _lib.OFB_F64


def blur(frames, sigma, out=None):
    """K1: Gaussian blur of ``frames [F,Nx,Ny]`` (mode nearest, truncate 4) -> new tensor."""
    _require_cuda(frames)
    frames = frames.contiguous()
    F, Nx, Ny = frames.shape
    wd, radius = _device_weights(sigma, frames.device)
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


def solve(P, b, x, speed_alpha, remodelling_alpha, rtol=1e-6, max_it=1000, options=None, workspace=None, recurrence=None):
    """K4/K5: solve A x = b in place (x: in = initial guess, out = solution).

    Returns ``(iterations int32[B], relres float64[B], converged int32[B])`` as device tensors; ``relres`` is the true
    residual |b - A x| / |b| of the returned x (reference :1150-1151).  ``recurrence`` (optional float64[B] device
    tensor) receives the BiCGStab recurrence residual the iteration stopped on."""
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
                                     _ptr(recurrence), B, Nx, Ny, _code(P), _stream_ptr(P.device)))
    return iters, relres, conv


_plan_cache = {}


def _direct_plan(Nx, Ny, device):
    """Nested-dissection plan of an Nx x Ny frame on `device`: (ctypes array of ofb_nd_level, tensors kept alive, n_levels)."""
    from . import ndplan
    key = (int(Nx), int(Ny), str(device))
    hit = _plan_cache.get(key)
    if hit is None:
        levels = ndplan.build(int(Nx), int(Ny))
        structs = (_lib.NdLevel * len(levels))()
        keep = []
        for k, lv in enumerate(levels):
            own, border, slot = (torch.from_numpy(np.ascontiguousarray(a)).to(device) for a in (lv.own, lv.border, lv.parent_slot))
            keep += [own, border, slot]
            structs[k] = _lib.NdLevel(lv.n_nodes, lv.n_own, lv.n_border, 0, own.data_ptr(), border.data_ptr(), slot.data_ptr())
        if len(_plan_cache) > 4:
            _plan_cache.clear()
        hit = _plan_cache[key] = (structs, keep, len(levels))
    return hit


def direct_solve_estimate(Nx, Ny):
    """Rough device bytes of a direct solve of an Nx x Ny pair without building the plan (measured: 1.25 / 5.9 / 27.5 GB
    at 256^2 / 512^2 / 1024^2: ~ pixels^1.12)."""
    return int(27.5e9 * (float(Nx) * float(Ny) / 1048576.0) ** 1.12)


def direct_solve_fits(Nx, Ny, device):
    """True if the fronts of a direct solve fit in 70 % of the device memory that is free right now."""
    try:
        free = torch.cuda.mem_get_info(device)[0]
    except Exception:
        return False
    return Nx * Ny <= (1 << 22) and direct_solve_estimate(Nx, Ny) * 1.05 < 0.7 * free


def direct_solve(P, b, speed_alpha, remodelling_alpha, workspace=None, rtol=1e-6):
    """Sparse direct solve of A x = b (``ofb_direct_solve``: multifrontal LU in fp64 on a nested dissection of the pixel
    grid), the counterpart of the reference's SuperLU branch (optical_flow.py:1146-1147).
    Returns ``(x [B,3,Nx,Ny] of P's dtype, relres float64[B], converged int32[B])``; relres is the true residual of x,
    converged the verdict of ``solve`` on it (rtol plus what storing x in P's dtype leaves)."""
    _require_cuda(P, b)
    B, Nx, Ny = P.shape
    assert b.shape == (B, 3, Nx, Ny) and b.dtype == P.dtype and P.is_contiguous() and b.is_contiguous()
    lib = _lib.load()
    structs, _, n_levels = _direct_plan(Nx, Ny, P.device)
    a, be = _weights(speed_alpha, B, P.device), _weights(remodelling_alpha, B, P.device)
    need = lib.ofb_direct_workspace_bytes(structs, n_levels, B, Nx, Ny)
    if need == 0:
        raise _lib.OfbError('direct solve: ' + lib.ofb_last_error().decode())
    holder = workspace if workspace is not None else Workspace()
    if holder.buffer is None or holder.buffer.numel() < need or holder.buffer.device != P.device:
        holder.buffer = None                                   # release the old one before asking for tens of GB
        holder.buffer = torch.empty(need, dtype=torch.uint8, device=P.device)
    x = torch.empty_like(b)
    relres = torch.zeros(B, dtype=torch.float64, device=P.device)
    conv = torch.zeros(B, dtype=torch.int32, device=P.device)
    _lib.check(lib.ofb_direct_solve(_ptr(P), _ptr(b), _ptr(x), _ptr(a), _ptr(be), structs, n_levels, _ptr(holder.buffer), need,
                                    float(rtol), _ptr(relres), _ptr(conv), B, Nx, Ny, _code(P), _stream_ptr(P.device)))
    return x, relres, conv


def last_unconverged():
    """Pairs the calling thread's last ``solve`` left unconverged (-1 if unknown): a host integer, no device read."""
    return int(_lib.load().ofb_last_unconverged())


def last_restarts():
    """Pairs the calling thread's last ``solve`` restarted from their true residual (a host integer)."""
    return int(_lib.load().ofb_last_restarts())


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


# ------------------------------------------------------------------------------------------------
# callers either side of the hot path (SURVEY.md 8f NEXT-3 / NEXT-4)
# ------------------------------------------------------------------------------------------------
def liu_shen(P, C, vx, vy, alpha, iterations, scale=1.0):
    """``iterations`` Liu-Shen Jacobi sweeps on ``B`` pairs (optical_flow.py:507-651); ``vx, vy [B,Nx,Ny]`` hold the initial
    guess in pixel units and are overwritten with the scaled velocities.  Returns ``(vx, vy, speed)``."""
    _require_cuda(P, C, vx, vy)
    B, Nx, Ny = P.shape
    for t in (C, vx, vy):
        assert t.is_contiguous() and t.shape == P.shape and t.dtype == P.dtype
    assert P.is_contiguous()
    speed, tx, ty = torch.empty_like(vx), torch.empty_like(vx), torch.empty_like(vy)
    _lib.check(_lib.load().ofb_liu_shen(_ptr(P), _ptr(C), _ptr(vx), _ptr(vy), _ptr(speed), _ptr(tx), _ptr(ty), float(alpha),
                                        int(iterations), float(scale), B, Nx, Ny, _code(P), _stream_ptr(P.device)))
    return vx, vy, speed


def movie_max(frames):
    """np.max of a device movie as a one-element float64 device tensor."""
    _require_cuda(frames)
    frames = frames.contiguous()
    out = torch.empty(1, dtype=torch.float64, device=frames.device)
    _lib.check(_lib.load().ofb_movie_max(_ptr(frames), frames.numel(), _code(frames), _ptr(out), _stream_ptr(frames.device)))
    return out


def adaptive_threshold(frames, window_size, threshold):
    """optical_flow.py:328-335 on device frames ``[F,Nx,Ny]`` -> bool tensor."""
    _require_cuda(frames)
    frames = frames.contiguous()
    F, Nx, Ny = frames.shape
    mx = movie_max(frames)
    tmp = torch.empty((3, F, Nx, Ny), dtype=torch.uint8, device=frames.device)      # 16-bit row sums + the uint8 plane
    out = torch.empty((F, Nx, Ny), dtype=torch.uint8, device=frames.device)
    _lib.check(_lib.load().ofb_adaptive_threshold(_ptr(frames), _ptr(mx), _ptr(tmp), _ptr(out), int(window_size),
                                                  int(np.ceil(threshold)), F, Nx, Ny, _code(frames), _stream_ptr(frames.device)))
    return out.view(torch.bool)


def clahe(frames, clip_limit, tiles_x, tiles_y, frames_in_flight=8):
    """optical_flow.py:362-370 on device frames ``[F,Nx,Ny]`` (values are truncated to uint16 like ``astype``)."""
    _require_cuda(frames)
    frames = frames.contiguous()
    F, Nx, Ny = frames.shape
    lib = _lib.load()
    in_flight = max(1, min(int(frames_in_flight), F))
    nbytes = lib.ofb_clahe_workspace_bytes(in_flight, int(tiles_x), int(tiles_y))
    if nbytes == 0:
        raise _lib.OfbError('CLAHE tile grid must be at least 1 x 1 (got %r x %r)' % (tiles_x, tiles_y))
    workspace = torch.empty(nbytes, dtype=torch.uint8, device=frames.device)
    out = torch.empty_like(frames)
    _lib.check(lib.ofb_clahe(_ptr(frames), _ptr(out), float(clip_limit), int(tiles_x), int(tiles_y), _ptr(workspace), nbytes,
                             F, Nx, Ny, _code(frames), _stream_ptr(frames.device)))
    return out


def subtract_background(frames, blurred, background):
    """optical_flow.py:195-198: ``frames - background`` where ``blurred > background``, else 0."""
    _require_cuda(frames, blurred)
    frames, blurred = frames.contiguous(), blurred.contiguous()
    assert frames.shape == blurred.shape and frames.dtype == blurred.dtype
    out = torch.empty_like(frames)
    _lib.check(_lib.load().ofb_subtract_background(_ptr(frames), _ptr(blurred), _ptr(out), float(background), frames.numel(),
                                                   _code(frames), _stream_ptr(frames.device)))
    return out


def widen(src, dst):
    """float32 device tensor -> float64 device tensor of the same shape (``ofb_widen``), on the current stream."""
    _require_cuda(src, dst)
    assert src.dtype == torch.float32 and dst.dtype == torch.float64 and src.shape == dst.shape
    assert src.is_contiguous() and dst.is_contiguous()
    _lib.check(_lib.load().ofb_widen(_ptr(src), _ptr(dst), src.numel(), _stream_ptr(src.device)))
    return dst
