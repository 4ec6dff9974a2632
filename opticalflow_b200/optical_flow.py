"""Drop-in for the hot path of kursawe/OpticalFlow ``source/optical_flow.py``.

Same entry points, argument conventions and result dictionaries as the reference
(``variational_optical_flow`` :715-1210, ``vary_regularisation`` :1918-1998, ``blur_movie`` :282-306, the
small helpers :676-713, :1212-1316, and the callers either side of the path: ``apply_adaptive_threshold``
:308-338, ``apply_clahe`` :340-374, the background subtraction :195-198, ``liu_shen_optical_flow_jit`` :426-673
with its driver ``conduct_variational_optical_flow_deprecated`` :1318-1529); the work is done by hand-written sm_100a kernels behind the C ABI of
``include/ofb200.h``.  There is no CPU fallback: without a CUDA device or without ``libofb200.so`` the
calls raise.

Differences from the reference that a caller can observe (all documented in DESIGN.md):
  * the frame pairs of a movie are solved as independent systems, batched, each from the user's
    ``initial_*`` guess (the reference warm-starts pair k from pair k-1, :803-806); the answers agree to
    O(rtol);
  * ``use_direct_solver=True`` runs a sparse direct solve on the device (multifrontal LU on a nested dissection
    of the pixel grid, fp64) instead of SuperLU; frames whose fronts do not fit in device memory (beyond
    ~2048^2) get the iterative solve with rtol 1e-12 instead.  The same direct solve takes over the pairs the
    iterative solver does not converge (small speed_alpha: the operator is indefinite there);
  * nothing is printed unless ``verbose=True``;
  * solver diagnostics of every pair are returned under ``result['solver_info']``.
Keyword-only extras (defaults preserve reference behaviour): ``dtype``, ``rtol``, ``max_it``,
``preconditioner``, ``device``, ``batch_pairs``, ``verbose``.
"""
import functools
import os
import time

import numpy as np
import torch

from . import _lib, engine, hostio, sharding

__all__ = ['variational_optical_flow', 'vary_regularisation', 'blur_movie', 'apply_numerical_derivative',
           'apply_constant_boundary_condition', 'get_index_set', 'format_elapsed_time', 'apply_adaptive_threshold',
           'apply_clahe', 'subtract_background', 'liu_shen_optical_flow_jit', 'conduct_variational_optical_flow_deprecated']

_QUANTITY_OFFSET = {'ux': 0, 'uy': 1, 'remodelling': 2}


# ------------------------------------------------------------------------------------------------
# small helpers of the reference interface (host-side index/array utilities)
# ------------------------------------------------------------------------------------------------
def format_elapsed_time(time_difference):
    """(minutes, seconds, milliseconds) of a time.time() difference (optical_flow.py:1212-1238)."""
    whole = int(time_difference)
    return int(time_difference // 60), int(time_difference % 60), int((time_difference - whole) * 1000)


def get_index_set(N_i, N_j, i_offset, j_offset, quantity, include_boundaries=False):
    """Flattened indices 3*N_j*(i+i_offset) + 3*(j+j_offset) + q of the reference's interleaved unknown
    numbering (optical_flow.py:1241-1302); q = 0,1,2 for 'ux','uy','remodelling'."""
    if quantity not in _QUANTITY_OFFSET:
        raise ValueError('I cannot figure out what quantity you want from me. Must be ux, uy, or remodelling.')
    lo = 0 if include_boundaries else 1
    i = np.arange(lo, N_i - lo, dtype=np.int32)[:, None] + i_offset
    j = np.arange(lo, N_j - lo, dtype=np.int32)[None, :] + j_offset
    return (3 * N_j * i + 3 * j + _QUANTITY_OFFSET[quantity]).ravel()


# rule -> list of (weight, row shift, column shift); shifts are relative to the interior pixel
_STENCILS = {
    'dx': ((0.5, 1, 0), (-0.5, -1, 0)),
    'dxy': ((0.25, 1, 1), (-0.25, 1, -1), (-0.25, -1, 1), (0.25, -1, -1)),
    'dxx': ((1.0, 1, 0), (1.0, -1, 0), (-2.0, 0, 0)),
    'dyy': ((1.0, 0, 1), (1.0, 0, -1), (-2.0, 0, 0)),
    'bar_x': ((1.0, 1, 0), (1.0, -1, 0)),
    'bar_y': ((1.0, 0, 1), (1.0, 0, -1)),
    'bar': ((1.0, -1, 0), (1.0, 1, 0), (1.0, 0, 1), (1.0, 0, -1)),
}
_STENCILS['dy'] = _STENCILS['dx']      # the reference's 'dy' rule is the x-derivative (:698-699); kept for parity
_STENCILS['dyx'] = _STENCILS['dxy']


def apply_numerical_derivative(matrix, rule):
    """Finite-difference stencil on a bordered 2-D array, interior returned (optical_flow.py:676-713)."""
    if rule not in _STENCILS:
        raise ValueError('unknown derivative rule %r' % (rule,))
    matrix = np.asarray(matrix)
    Nx, Ny = matrix.shape
    out = np.zeros((Nx - 2, Ny - 2), dtype=np.result_type(matrix.dtype, np.float64))
    for weight, di, dj in _STENCILS[rule]:
        out += weight * matrix[1 + di:Nx - 1 + di, 1 + dj:Ny - 1 + dj]
    return out


def apply_constant_boundary_condition(image):
    """In-place edge overwrite, rows first then columns (optical_flow.py:1304-1316)."""
    for axis in (0, 1):
        view = image if axis == 0 else image.T
        view[0] = view[2]
        view[-1] = view[-3]


def _device(device):
    if not torch.cuda.is_available():
        raise _lib.OfbError('no CUDA device available: this path has no CPU fallback')
    return torch.device('cuda', torch.cuda.current_device()) if device is None else torch.device(device)


def _on_device(fn):
    """Run ``fn`` with the requested GPU as the calling thread's current CUDA device.  ``libofb200.so`` keys its per-device
    state (SM count, occupancy, side stream, poll words, function attributes) on ``cudaGetDevice()`` and launches on the
    stream torch made for that device; a fresh host thread - the ``devices=[...]`` driver starts one per GPU - has device 0
    current, so every entry point sets the device itself."""
    @functools.wraps(fn)
    def wrapper(*args, **kwargs):
        device = kwargs.get('device')
        devices = kwargs.get('devices')
        if device is None and devices is not None and len(devices) == 1:
            device = devices[0]
        with torch.cuda.device(_device(device)):
            return fn(*args, **kwargs)
    return wrapper


@_on_device
def blur_movie(movie, smoothing_sigma, *, dtype='float64', device=None):
    """Per-frame Gaussian blur (optical_flow.py:282-306), computed on the GPU, returned as float64."""
    dev = _device(device)
    tdtype, _ = engine.resolve_dtype(dtype)
    movie = np.asarray(movie)
    frames = torch.from_numpy(np.ascontiguousarray(movie)).to(dev).to(tdtype)
    return engine.blur(frames, smoothing_sigma).to(torch.float64).cpu().numpy()


# ------------------------------------------------------------------------------------------------
# the hot path
# ------------------------------------------------------------------------------------------------
_copy_streams = {}


def _copy_stream(dev):
    key = str(dev)
    if key not in _copy_streams:
        _copy_streams[key] = torch.cuda.Stream(dev)
    return _copy_streams[key]


def _default_batch(Nx, Ny, itemsize, n_pairs, dev=None, pipelined=True):
    """Frame pairs per chunk.  Larger chunks are faster per pair - 16 / 32 / 64 pairs of 1024^2 solve at 775 / 887 / 943
    pairs per second on a B200: the persistent sweeps get longer row ranges per CTA and the one-CTA-per-pair coarse
    kernels more CTAs - until ~64.  ~16 level-0 vectors of workspace per pair; bounded by a quarter of the free device
    memory and 16 GiB."""
    per_pair = 16 * 3 * Nx * Ny * itemsize
    budget = 16 << 30
    if dev is not None:
        try:
            budget = min(budget, torch.cuda.mem_get_info(dev)[0] // 4)
        except Exception:
            pass
    B = int(max(1, min(n_pairs, budget // per_pair, 64)))
    if not pipelined:                              # nothing is copied back per chunk (vary_regularisation): just fill the device
        return B
    # The egress of a chunk (device->host copy, float64 widening: ~0.7 ms per 1024^2 pair on 8 host threads) only starts
    # when the chunk is solved, so a movie of few chunks is bound by first solve + whole egress, not by the device: keep at
    # least ~4 chunks in a call as long as a chunk still has >= 16 Mpixel (128 pairs: 32 per chunk, 500 pairs: 64).
    # With direct egress (no host stage) the same holds for the copy engine's backlog: 128 pairs as 32-32-32-16-16 take
    # 180 ms, as 64-48-16 190 ms (the float64 planes of the 48-pair chunk are still on the wire when the last solve ends).
    floor16 = -(-(16 << 20) // (Nx * Ny))
    B = min(B, max(floor16, n_pairs // 4))
    # one chunk would leave nothing to overlap the device->host egress with: split in two when both halves
    # still fill the machine (>= 4 Mpixel per launch)
    if B >= n_pairs and n_pairs * Nx * Ny >= (8 << 20):
        B = max((n_pairs + 1) // 2, -(-(4 << 20) // (Nx * Ny)))
    return max(1, min(B, n_pairs))


def _chunk_schedule(n_pairs, B, pixels, taper=True):
    """Pair ranges [(k0, k1), ...] of the chunk pipeline: chunks of B and, with ``taper``, one short final chunk of
    ~B/4 pairs (at least ~16 Mpixel: below that the fixed costs of a solve dominate - 4-pair chunks of 1024^2 run at
    2.0 ms per pair instead of 1.4).  Nothing overlaps the device->host copy and the float64 widening of the final
    chunk, so its size sets the tail of the call: a 64-pair chunk of 1024^2 leaves ~58 ms, a 16-pair one 15 ms."""
    plain = [(k, min(k + B, n_pairs)) for k in range(0, n_pairs, B)]
    last = max(-(-(16 << 20) // max(1, pixels)), B // 4)
    if not taper or n_pairs <= B or last >= B or n_pairs - last < 1:
        return plain
    main = n_pairs - last
    n_chunks = -(-main // B)
    bounds, k = [], 0
    for c in range(n_chunks):                       # the main part in chunks of B, the short one of them next to the end
        size = B if (c + 1) * B <= main else main - c * B
        bounds.append((k, k + size))
        k += size
    bounds.append((k, n_pairs))
    return bounds


class _EventWaiter:
    """``wait()`` on a CUDA event, the interface the staged path's worker futures have."""

    def __init__(self, event):
        self.event = event

    def wait(self):
        self.event.synchronize()


PROBE_IT = 100


def _solve_pairs(P, C, alpha, beta, x0, rtol, max_it, preconditioner, workspace, solver_options, direct=False, direct_workspace=None):
    """Solve a batch.  ``direct``: every pair by the sparse direct solve (the reference's ``use_direct_solver=True`` branch,
    :1146-1147).  Otherwise multigrid-preconditioned BiCGStab; pairs it has not converged after PROBE_IT iterations are in
    the regime where the operator is indefinite (speed_alpha below ~max I^2 / 1.5) and go to the direct
    solve when its fronts fit on the device (0.22 s per 1024^2 pair).  Where they do not fit (very large frames) the older chain
    applies: the robust V(2,2) cycle, float64, block-Jacobi; a pair none of them converges is reported like the reference
    reports it (:1137-1138): ``converged`` False, best iterate returned."""
    B, Nx, Ny = P.shape
    b = engine.build_rhs(P, C)
    fallback = torch.zeros(B, dtype=torch.bool, device=P.device)
    direct_workspace = direct_workspace if direct_workspace is not None else engine.Workspace()
    if direct:
        x, relres, conv = engine.direct_solve(P, b, alpha, beta, direct_workspace, rtol)
        return x, torch.zeros(B, dtype=torch.int32, device=P.device), relres, conv, fallback, relres.clone()
    x = x0.clone()
    opts = engine.make_options(preconditioner, **solver_options)
    recur = torch.zeros(B, dtype=torch.float64, device=P.device)
    if opts.precond != _lib.PRECOND_MG:
        iters, relres, conv = engine.solve(P, b, x, alpha, beta, rtol, max_it, opts, workspace, recurrence=recur)
        return x, iters, relres, conv, fallback, recur
    # The cycle needs 6-10 iterations where it works; a pair still running after PROBE_IT iterations is in a regime
    # it does not handle, and the remaining budget is better spent on the fallbacks (which restart from x0 anyway).
    probe = min(int(max_it), PROBE_IT)
    iters, relres, conv = engine.solve(P, b, x, alpha, beta, rtol, probe, opts, workspace, recurrence=recur)
    if engine.last_unconverged() == 0:
        # every pair converged: no fallback, and no device->host read either - a read here would queue on the copy
        # engine behind the previous chunk's result planes (tens of ms for a 64-pair chunk)
        return x, iters, relres, conv, fallback, recur
    a_all, be_all = engine._weights(alpha, B, P.device), engine._weights(beta, B, P.device)
    if os.environ.get('OFB_DIRECT_FALLBACK', '1') != '0' and engine.direct_solve_fits(Nx, Ny, P.device):
        bad = (conv == 0).nonzero().flatten()
        xb, rr2, cv2 = engine.direct_solve(P[bad].contiguous(), b[bad].contiguous(), a_all[bad], be_all[bad], direct_workspace, rtol)
        x[bad] = xb
        relres[bad] = rr2
        recur[bad] = rr2
        conv[bad] = cv2
        fallback[bad] = True
        return x, iters, relres, conv, fallback, recur
    passthrough = {k: v for k, v in solver_options.items() if k in ('check_every', 'rows_per_strip', 'flags')}
    robust = engine.make_options('mg', nu_pre=2, nu_post=2, omega=0.8, omega_pre=(0.8, 0.8), omega_post=(0.8, 0.8), **passthrough)
    chain = []                                      # (options, solve in float64, iteration budget)
    custom = any(k in solver_options for k in ('nu_pre', 'nu_post', 'omega', 'omega_pre', 'omega_post'))
    if P.dtype == torch.float32:
        if not custom:
            chain.append((robust, False, probe))
        chain.append((robust, True, max_it))
    else:
        chain.append((opts if custom else robust, False, max_it))
    chain.append((engine.make_options('bjacobi', **passthrough), P.dtype == torch.float32, max_it))
    for fb_opts, widen, budget in chain:
        bad = (conv == 0).nonzero().flatten()
        if bad.numel() == 0:
            break
        a_b, be_b = a_all[bad], be_all[bad]
        if widen:
            Pb, Cb = P[bad].double().contiguous(), C[bad].double().contiguous()
            bb, xb, ws = engine.build_rhs(Pb, Cb), x0[bad].double().contiguous(), engine.Workspace()
        else:
            Pb, bb, xb, ws = P[bad].contiguous(), b[bad].contiguous(), x0[bad].clone(), workspace
        rc2 = torch.zeros(bad.numel(), dtype=torch.float64, device=P.device)
        it2, rr2, cv2 = engine.solve(Pb, bb, xb, a_b, be_b, rtol, budget, fb_opts, ws, recurrence=rc2)
        better = (cv2 == 1) | (rr2 < relres[bad])
        sel = bad[better]
        x[sel] = xb[better].to(x.dtype)
        iters[sel] = it2[better] + iters[sel]
        relres[sel] = rr2[better]
        recur[sel] = rc2[better]
        conv[sel] = cv2[better]
        fallback[sel] = True
    return x, iters, relres, conv, fallback, recur


@_on_device
def variational_optical_flow(movie, delta_x=1.0, delta_t=1.0, speed_alpha=1.0, remodelling_alpha=1000.0,
                             smoothing_sigma=None, initial_v_x=0.0, initial_v_y=0.0, initial_remodelling=0.0,
                             use_direct_solver=False, *, dtype='float64', rtol=1e-6, max_it=1000,
                             preconditioner='mg', device=None, batch_pairs=None, verbose=False,
                             solver_options=None, devices=None, result_dtype='float64', outputs=None, warm_start=False):
    """Variational optical flow with net remodelling on every consecutive frame pair of ``movie``.

    Signature, defaults, units and result keys follow optical_flow.py:715-1210: ``movie[T,Nx,Ny]``; velocities
    are returned in ``delta_x/delta_t`` units, ``remodelling`` unscaled; ``result['converged']`` refers to the
    last pair (:1202) and ``result['speed_functional']`` holds the remodelling functional (:1205).

    The frame pairs are processed in chunks of ``batch_pairs``: while the device solves chunk k+1, chunk k's
    planes travel device -> pinned memory on a copy stream and are widened to float64 by a worker thread.
    ``devices=[...]`` shards the pairs over several GPUs (contiguous ranges, one host thread per device, no
    collective; sharding.py).

    ``warm_start=True`` follows the reference's habit of seeding a pair with an earlier pair's solution (:803-806): the
    pairs of a chunk are solved together, so every pair of chunk c starts from the solution of the LAST pair of chunk
    c-1 (the first chunk from ``initial_*``).  The answer changes within the solver tolerance only; on slowly varying
    flow it saves iterations.  Off by default: pairs then depend on the chunking.

    The reference returns ~42 MB of float64 per 1024^2 pair (:1189-1205); at several hundred pairs per second and GPU
    that is what the host link carries, not the solve.  Two opt-in keywords cut it (defaults keep the reference's
    dictionary): ``result_dtype='float32'`` returns the field arrays, ``blurred_data`` and ``original_data`` as
    float32 (``original_data`` is then the caller's movie itself when that already is float32);
    ``outputs=('v_x', 'v_y', 'remodelling')`` names the arrays to ship - keys left out of
    ``('v_x', 'v_y', 'speed', 'remodelling', 'blurred_data', 'original_data')`` are absent from the result
    (``speed`` is ``np.hypot(v_x, v_y)``, ``blurred_data`` is ``blur_movie(movie, smoothing_sigma)``)."""
    if devices is not None and len(devices) > 1:
        kw = dict(delta_x=delta_x, delta_t=delta_t, speed_alpha=speed_alpha, remodelling_alpha=remodelling_alpha,
                  smoothing_sigma=smoothing_sigma, initial_v_x=initial_v_x, initial_v_y=initial_v_y,
                  initial_remodelling=initial_remodelling, use_direct_solver=use_direct_solver, dtype=dtype, rtol=rtol,
                  max_it=max_it, preconditioner=preconditioner, batch_pairs=batch_pairs, verbose=verbose,
                  solver_options=solver_options, result_dtype=result_dtype, outputs=outputs, warm_start=warm_start)
        return sharding.solve_on_devices(movie, list(devices), lambda frames, dev: variational_optical_flow(frames, device=dev, **kw),
                                         delta_x, delta_t)
    if devices is not None and len(devices) == 1:
        device = devices[0]
    t_start = time.time()
    dev = _device(device)
    tdtype, _ = engine.resolve_dtype(dtype)
    movie_np = np.asarray(movie)
    if movie_np.ndim != 3 or movie_np.shape[0] < 2:
        raise ValueError('movie must be [T>=2, Nx, Ny]')
    T, Nx, Ny = movie_np.shape
    n_pairs = T - 1
    # use_direct_solver=True (:1146-1147): the sparse direct solve on the device when its fronts fit (frames up to ~2048^2 on a
    # B200), else the iterative solve with the tightest tolerance
    direct_mode = bool(use_direct_solver) and engine.direct_solve_fits(Nx, Ny, dev)
    if use_direct_solver and not direct_mode:
        rtol = min(rtol, 1e-12)
    solver_options = dict(solver_options or {})
    itemsize = 4 if tdtype == torch.float32 else 8
    blur = smoothing_sigma is not None
    arena = hostio.result_arena                                                # recycles dropped result arrays' pages
    wire = np.dtype(result_dtype)
    if wire not in (np.dtype(np.float32), np.dtype(np.float64)):
        raise ValueError("result_dtype must be 'float64' or 'float32', got %r" % (result_dtype,))
    wire_t = torch.float32 if wire == np.float32 else torch.float64
    all_outputs = ('v_x', 'v_y', 'speed', 'remodelling', 'blurred_data', 'original_data')
    wanted = all_outputs if outputs is None else tuple(outputs)
    if set(wanted) - set(all_outputs):
        raise ValueError('outputs must be taken from %r, got %r' % (all_outputs, outputs))
    result_names = ('v_x', 'v_y', 'speed', 'remodelling')
    shipped = tuple(q for q, name in enumerate(result_names) if name in wanted)     # plane numbers that leave the device
    ship_blur = blur and 'blurred_data' in wanted
    # OFB_HOST_SPEED=1: speed = sqrt(v_x^2 + v_y^2) (:1191) computed on the host from the two arrays that arrive anyway
    # instead of being shipped (8 of the 42 MB per 1024^2 pair less on the host link).  Off by default: measured on the
    # pool's 8-GPU box it gains nothing at 1-2 ranks (789 / 1169 vs 1147 frame-pairs/s) and loses at 4 and 8 (1228 vs
    # 1277, 1414 vs 1713), where the host's memory system, not the link, is what saturates (profiles/r02_d2h_ceiling.txt)
    host_speed = all(k in wanted for k in ('v_x', 'v_y', 'speed')) and os.environ.get('OFB_HOST_SPEED', '0') == '1'
    result_bytes = ((len(shipped) - (1 if host_speed else 0)) * n_pairs + (T if ship_blur else 0)) * Nx * Ny * wire.itemsize
    direct = (os.environ.get('OFB_DIRECT_EGRESS', '1') != '0' and arena.max_idle_bytes > 0
              and result_bytes <= hostio.direct_egress_limit())
    B = int(batch_pairs) if batch_pairs else _default_batch(Nx, Ny, itemsize, n_pairs, dev)
    B = max(1, min(B, n_pairs))
    scale_in = float(delta_t) / float(delta_x)                                 # :800-801
    scale_out = float(delta_x) / float(delta_t)                                # :1189-1190
    guess = (float(initial_v_x) * scale_in, float(initial_v_y) * scale_in, float(initial_remodelling))

    source = np.ascontiguousarray(movie_np)
    if 'original_data' not in wanted:
        original = None
    elif source.dtype == wire:
        original = source                                                      # nothing to convert (float32 contract only)
    else:
        original = arena.alloc((T, Nx, Ny), wire)                              # :769
    convert_original = original is not None and original is not source
    # Direct egress: the result arrays are page-locked, the planes are converted to the result type on the device and
    # copied straight into them - the host cores do no conversion (on a slow host the float32 -> float64 widening, not
    # the device, bounded the call).  Pinning costs ~0.1-0.3 s per GB the first time (recycled by the arena afterwards);
    # it is used up to hostio.direct_egress_limit() bytes of results per call; beyond that, or if pinning fails, the
    # planes go through pinned staging and the host copy pool.
    out, blurred, out_t, blurred_t = None, None, None, None
    if direct:
        try:
            out = {result_names[q]: arena.alloc((n_pairs, Nx, Ny), wire, pinned=not (host_speed and q == 2)) for q in shipped}
            blurred = arena.alloc((T, Nx, Ny), wire, pinned=True) if ship_blur else None
            out_t = {k: torch.from_numpy(v) for k, v in out.items() if not (host_speed and k == 'speed')}
            blurred_t = torch.from_numpy(blurred) if ship_blur else None
            direct = all(t.is_pinned() for t in out_t.values()) and (blurred_t is None or blurred_t.is_pinned())
        except Exception:
            direct = False
    if not direct:
        host_speed = False                                                      # the staged path's host threads are busy widening
        out = {result_names[q]: arena.alloc((n_pairs, Nx, Ny), wire) for q in shipped}
        blurred = arena.alloc((T, Nx, Ny), wire) if ship_blur else None
    wire_planes = tuple(q for q in shipped if not (host_speed and q == 2))        # planes copied device -> host
    functionals = np.zeros((n_pairs, 3))
    # relative_residual: |b - A x| / |b| recomputed from the returned x (what the reference prints, :1150-1151);
    # recurrence_residual: the solver's own norm when it stopped (:1140-1142)
    info = dict(iterations=np.zeros(n_pairs, np.int32), relative_residual=np.zeros(n_pairs),
                recurrence_residual=np.zeros(n_pairs), converged=np.zeros(n_pairs, bool), fallback=np.zeros(n_pairs, bool))
    workspace = engine.Workspace()
    direct_workspace = engine.Workspace()                                      # allocated only if a direct solve happens
    nplanes = 5 if blur else 4                                                 # v_x, v_y, speed, remodelling (, blurred)
    main_stream = torch.cuda.current_stream(dev)
    copy_stream = _copy_stream(dev)
    worker = hostio.EgressWorker()
    uploader = hostio.FrameUploader(B + 1, (Nx, Ny), source.dtype, dev)
    slots = []                                                                 # device + pinned planes, three deep: chunk k+2 is
    n_slots = 3 if n_pairs > 2 * B else 2                                      # solved while k+1 is copied and k is widened
    for s in range(n_slots):
        slot = dict(dev=torch.empty((nplanes, B + 1, Nx, Ny), dtype=tdtype, device=dev), pin=None, busy=None)
        if not direct:
            slot['pin'] = hostio.pinned_pool.acquire((nplanes, B + 1, Nx, Ny), tdtype)
        # planes in the result type for the direct path (planes that already have it are copied as they are)
        slot['dev64'] = (torch.empty((nplanes, B + 1, Nx, Ny), dtype=wire_t, device=dev)
                         if direct and tdtype != wire_t else slot['dev'])
        slots.append(slot)
    timeline = []
    host_jobs = []                                                             # host-side work on result arrays (speed, input copy)
    previous_last = None                                                       # warm_start: solution of the previous chunk's last pair
    t_setup = time.time()
    try:
        schedule = _chunk_schedule(n_pairs, B, Nx * Ny, taper=os.environ.get('OFB_TAPER', '1') != '0')
        uploader.prefetch(source[schedule[0][0]:schedule[0][1] + 1])
        stats = []
        for c, (k0, k1) in enumerate(schedule):
            nb = k1 - k0
            slot = slots[c % n_slots]
            t_chunk = time.perf_counter()
            if slot['busy'] is not None:
                slot['busy'].wait()                                            # worker has drained this slot's pinned planes
            frames = uploader.take(tdtype)                                     # nb+1 frames, uploaded during the previous solve
            t_take = time.perf_counter()
            if c + 1 < len(schedule):
                uploader.prefetch(source[schedule[c + 1][0]:schedule[c + 1][1] + 1])
            if blur:
                frames = engine.blur(frames, smoothing_sigma, out=slot['dev'][4, :nb + 1])     # :770-771
            P, C = frames[:-1], frames[1:]
            x0 = torch.empty((nb, 3, Nx, Ny), dtype=tdtype, device=dev)
            if warm_start and previous_last is not None:
                x0[:] = previous_last                                          # :803-806, across chunks
            else:
                for q in range(3):
                    x0[:, q] = guess[q]
            t_prep = time.perf_counter()
            x, iters, relres, conv, fb, recur = _solve_pairs(P, C, speed_alpha, remodelling_alpha, x0, rtol, max_it,
                                                             preconditioner, workspace, solver_options, direct_mode, direct_workspace)
            t_solve = time.perf_counter()
            if warm_start:
                previous_last = x[-1:].clone()
            fields = slot['dev'][:4, :nb]
            _, _, _, _, fun, _ = engine.epilogue(P, C, x, speed_alpha, remodelling_alpha, scale_out, out=fields)
            last_frame = nb + 1 if k1 == n_pairs else nb
            if direct:
                planes = slot['dev64']
                if planes is not slot['dev']:
                    convert = engine.widen if wire_t == torch.float64 else (lambda src, dst: dst.copy_(src))
                    for q in wire_planes:
                        convert(slot['dev'][q, :nb], planes[q, :nb])
                    if ship_blur:
                        convert(slot['dev'][4, :last_frame], planes[4, :last_frame])
                ready = torch.cuda.Event()
                ready.record(main_stream)
                with torch.cuda.stream(copy_stream):
                    copy_stream.wait_event(ready)
                    for q in wire_planes:
                        out_t[result_names[q]][k0:k1].copy_(planes[q, :nb], non_blocking=True)
                        if host_speed and q == 1:              # v_x and v_y of the chunk have landed: speed can start
                            landed = torch.cuda.Event()
                            landed.record(copy_stream)
                    if ship_blur:
                        blurred_t[k0:k0 + last_frame].copy_(planes[4, :last_frame], non_blocking=True)
                    copied = torch.cuda.Event()
                    copied.record(copy_stream)
                slot['busy'] = _EventWaiter(copied)            # slot reuse waits for the copy engine, not for a host thread
                if convert_original:
                    host_jobs.append(worker.submit(None, [(original[k0:k0 + last_frame], source[k0:k0 + last_frame])]))   # :769
                if host_speed:
                    host_jobs.append(worker.submit(landed, [], functools.partial(hostio.speed_from_velocities, out['v_x'][k0:k1],
                                                                                 out['v_y'][k0:k1], out['speed'][k0:k1])))
            else:
                ready = torch.cuda.Event()
                ready.record(main_stream)
                with torch.cuda.stream(copy_stream):
                    copy_stream.wait_event(ready)
                    for q in shipped:
                        slot['pin'][q, :nb].copy_(slot['dev'][q, :nb], non_blocking=True)
                    if ship_blur:
                        slot['pin'][4, :last_frame].copy_(slot['dev'][4, :last_frame], non_blocking=True)
                    copied = torch.cuda.Event()
                    copied.record(copy_stream)
                # one widening job per chunk: splitting it per plane was measured slower (27 vs 22 ms per chunk)
                copies = [(out[result_names[q]][k0:k1], slot['pin'][q, :nb]) for q in shipped]
                if ship_blur:
                    copies.append((blurred[k0:k0 + last_frame], slot['pin'][4, :last_frame]))
                if convert_original:
                    copies.append((original[k0:k0 + last_frame], source[k0:k0 + last_frame]))      # float64 copy of the input (:769)
                slot['busy'] = worker.submit(copied, copies)       # slot reuse (three chunks later) waits for this on the host
            stats.append(torch.cat([fun, iters.double()[:, None], relres[:, None], conv.double()[:, None],
                                    fb.double()[:, None], recur[:, None]], dim=1))
            timeline.append(dict(chunk=c, wait_slot_and_frames=t_take - t_chunk, prepare=t_prep - t_take,
                                 solve=t_solve - t_prep, submit=time.perf_counter() - t_solve))
        t_loop = time.perf_counter()
        packed = torch.cat(stats, dim=0).cpu().numpy()                         # one device->host read for all diagnostics
        functionals[:] = packed[:, 0:3]
        info['iterations'][:] = packed[:, 3].astype(np.int32)
        info['relative_residual'][:] = packed[:, 4]
        info['converged'][:] = packed[:, 5] != 0
        info['fallback'][:] = packed[:, 6] != 0
        info['recurrence_residual'][:] = packed[:, 7]
        if verbose:
            print('iterations %s, residuals %s' % (info['iterations'], info['relative_residual']))
        t_stats = time.perf_counter()
        for slot in slots:
            if slot['busy'] is not None:
                slot['busy'].wait()
        for job in host_jobs:
            job.wait()
        timeline.append(dict(stats=t_stats - t_loop, drain=time.perf_counter() - t_stats, setup=t_setup - t_start))
    finally:
        # nothing may still be writing into the result arrays or the staging slots when they are handed back; an egress
        # worker error (re-raised by close()) comes last so that it cannot skip the other steps
        try:
            torch.cuda.current_stream(dev).synchronize()
            copy_stream.synchronize()
        finally:
            try:
                uploader.close()
            finally:
                try:
                    for slot in slots:
                        if slot['pin'] is not None:
                            hostio.pinned_pool.release(slot['pin'])
                finally:
                    worker.close()
    if verbose and not info['converged'].all():
        print('the solver has not actually converged, the result will be incorrect or inaccurate')
    result = dict()
    for name in result_names:
        if name in out:
            result[name] = out[name]
    if original is not None:
        result['original_data'] = original
    result['delta_x'] = delta_x
    result['delta_t'] = delta_t
    if 'blurred_data' in wanted:
        result['blurred_data'] = blurred if blur else original if original is not None else source.astype(wire)
    result['converged'] = bool(info['converged'][-1])
    result['L1_functional'] = float(np.sum(functionals[:, 0]))
    result['remodelling_functional'] = float(np.sum(functionals[:, 2]))
    result['speed_functional'] = float(np.sum(functionals[:, 2]))          # sic, optical_flow.py:1205
    info['speed_functional_per_pair'] = functionals[:, 1].copy()
    info['elapsed_seconds'] = time.time() - t_start
    info['dtype'] = str(tdtype).replace('torch.', '')
    info['batch_pairs'] = B
    info['timeline'] = timeline
    info['egress_log'] = list(worker.log)
    info['h2d_bytes'] = int(sum(k1 - k0 + 1 for k0, k1 in schedule) * source[0].nbytes)
    info['d2h_bytes'] = int((len(wire_planes) * n_pairs + (T if ship_blur else 0)) * Nx * Ny * (wire.itemsize if direct else itemsize))
    info['speed_on_host'] = bool(host_speed)
    info['direct_egress'] = bool(direct)
    info['direct_solver'] = bool(direct_mode)
    result['solver_info'] = info
    return result


def _sweep_pairs(movie_np, first_pair, last_pair, cell_alpha, cell_beta, dev, tdtype, smoothing_sigma, guess, scale, rtol, max_it,
                 preconditioner, solver_options, batch_entries, direct_mode=False):
    """Pairs [first_pair, last_pair) of the movie x every (alpha, beta) cell on one device.  Returns host arrays
    (moments [cells, 4], functional [cells], converged of the last pair handled [cells])."""
    n_cells = cell_alpha.size
    _, Nx, Ny = movie_np.shape
    a_dev = torch.from_numpy(cell_alpha).to(dev)
    b_dev = torch.from_numpy(cell_beta).to(dev)
    mom_dev = torch.zeros((n_cells, 4), dtype=torch.float64, device=dev)
    fun_dev = torch.zeros(n_cells, dtype=torch.float64, device=dev)
    conv_dev = torch.zeros(n_cells, dtype=torch.int32, device=dev)
    workspace = engine.Workspace()
    direct_workspace = engine.Workspace()
    n_here = last_pair - first_pair
    if n_here >= max(8, batch_entries // 2):
        # many pairs: a batch is `batch_entries` consecutive pairs of ONE cell.  Every entry of a batch then needs about the same
        # number of iterations - in a batch that mixes cells all entries wait for the slowest weight (6 iterations at alpha 1e5, 11
        # at 1e8), paying the fixed costs of the extra iterations - and the blurred frames are shared, not replicated per cell.
        for k0 in range(first_pair, last_pair, batch_entries):
            k1 = min(k0 + batch_entries, last_pair)
            npairs = k1 - k0
            frames = hostio.as_native_tensor(movie_np[k0:k1 + 1]).to(dev).to(tdtype)
            if smoothing_sigma is not None:
                frames = engine.blur(frames, smoothing_sigma)
            P, C = frames[:-1], frames[1:]
            x0 = torch.empty((npairs, 3, Nx, Ny), dtype=tdtype, device=dev)
            for q in range(3):
                x0[:, q] = guess[q]
            for c in range(n_cells):
                a, be = a_dev[c:c + 1].expand(npairs).contiguous(), b_dev[c:c + 1].expand(npairs).contiguous()
                x, iters, relres, conv, fb, _ = _solve_pairs(P, C, a, be, x0, rtol, max_it, preconditioner, workspace, solver_options,
                                                             direct_mode, direct_workspace)
                _, _, _, _, fun, mom = engine.epilogue(P, C, x, a, be, scale)
                mom_dev[c] += mom.sum(dim=0)
                fun_dev[c] += (fun[:, 0] + 2.0 * fun[:, 2]).sum()
                conv_dev[c] = conv[-1]
        packed = torch.cat([mom_dev, fun_dev[:, None], conv_dev.double()[:, None]], dim=1).cpu().numpy()
        return packed[:, :4], packed[:, 4], packed[:, 5] != 0
    # few pairs: entries = (pair, cell), pairs outermost, solved in batches that span pairs: 4 pairs x 16 cells fill a 64-entry solve
    pairs_per_batch = max(1, int(round(batch_entries / float(n_cells))))     # 11 cells: 6 pairs = 66 entries, not 5 = 55
    cells_per_batch = min(n_cells, batch_entries)
    for k0 in range(first_pair, last_pair, pairs_per_batch):
        k1 = min(k0 + pairs_per_batch, last_pair)
        frames = hostio.as_native_tensor(movie_np[k0:k1 + 1]).to(dev).to(tdtype)
        if smoothing_sigma is not None:
            frames = engine.blur(frames, smoothing_sigma)               # shared by every cell of the pair
        for c0 in range(0, n_cells, cells_per_batch):
            c1 = min(c0 + cells_per_batch, n_cells)
            nc, npairs = c1 - c0, k1 - k0
            # entry e = (pair k0 + e // nc, cell c0 + e % nc)
            P = frames[:-1].repeat_interleave(nc, dim=0)
            C = frames[1:].repeat_interleave(nc, dim=0)
            a = a_dev[c0:c1].repeat(npairs)
            be = b_dev[c0:c1].repeat(npairs)
            x0 = torch.empty((nc * npairs, 3, Nx, Ny), dtype=tdtype, device=dev)
            for q in range(3):
                x0[:, q] = guess[q]
            x, iters, relres, conv, fb, _ = _solve_pairs(P, C, a, be, x0, rtol, max_it, preconditioner, workspace, solver_options,
                                                         direct_mode, direct_workspace)
            _, _, _, _, fun, mom = engine.epilogue(P, C, x, a, be, scale)
            mom_dev[c0:c1] += mom.view(npairs, nc, 4).sum(dim=0)
            # L1 + 'speed' (= remodelling, :1205) + remodelling
            fun_dev[c0:c1] += (fun[:, 0] + 2.0 * fun[:, 2]).view(npairs, nc).sum(dim=0)
            conv_dev[c0:c1] = conv.view(npairs, nc)[-1]
    packed = torch.cat([mom_dev, fun_dev[:, None], conv_dev.double()[:, None]], dim=1).cpu().numpy()     # the one device -> host read
    return packed[:, :4], packed[:, 4], packed[:, 5] != 0


def vary_regularisation(movie, speed_alpha_values=np.arange(500, 2000, 500),
                        remodelling_alpha_values=np.arange(500, 2000, 500), filename=None, **kwargs):
    """Grid of regularisation weights -> summary statistics (optical_flow.py:1918-1998).

    Every (speed_alpha, remodelling_alpha) cell is what ``variational_optical_flow(movie, speed_alpha=...,
    remodelling_alpha=..., **kwargs)`` would give, but the cells are solved together: the blurred frames are
    shared on the device, the (pair, cell) entries form batches with per-entry weights that span frame pairs (4 pairs
    x 16 cells fill a 64-entry solve); means, variances and the functional are accumulated on the device (epilogue
    moments) and read once at the end - no field is copied back.  ``devices=[...]`` shards the frame pairs over several
    GPUs (pairs outermost, one host thread per device, partial sums added on the host)."""
    devices = kwargs.pop('devices', None)
    device = kwargs.pop('device', None)
    if devices is None or len(devices) == 0:
        devices = [device]
    devs = [_device(d) for d in devices]
    tdtype, _ = engine.resolve_dtype(kwargs.pop('dtype', 'float64'))
    rtol = kwargs.pop('rtol', 1e-6)
    max_it = kwargs.pop('max_it', 1000)
    preconditioner = kwargs.pop('preconditioner', 'mg')
    solver_options = dict(kwargs.pop('solver_options', None) or {})
    batch_pairs = kwargs.pop('batch_pairs', None)
    kwargs.pop('verbose', None)
    delta_x = kwargs.pop('delta_x', 1.0)
    delta_t = kwargs.pop('delta_t', 1.0)
    smoothing_sigma = kwargs.pop('smoothing_sigma', None)
    guess = (float(kwargs.pop('initial_v_x', 0.0)) * delta_t / delta_x,
             float(kwargs.pop('initial_v_y', 0.0)) * delta_t / delta_x, float(kwargs.pop('initial_remodelling', 0.0)))
    use_direct = bool(kwargs.pop('use_direct_solver', False))
    if kwargs:
        raise TypeError('unexpected keyword arguments: %s' % sorted(kwargs))
    movie_np = np.ascontiguousarray(np.asarray(movie))
    T, Nx, Ny = movie_np.shape
    direct_mode = use_direct and engine.direct_solve_fits(Nx, Ny, devs[0])
    if use_direct and not direct_mode:
        rtol = min(rtol, 1e-12)
    n_pairs = T - 1
    alphas = np.asarray(speed_alpha_values, dtype=np.float64)
    betas = np.asarray(remodelling_alpha_values, dtype=np.float64)
    cell_alpha = np.repeat(alphas, len(betas))
    cell_beta = np.tile(betas, len(alphas))
    n_cells = cell_alpha.size
    itemsize = 4 if tdtype == torch.float32 else 8
    scale = float(delta_x) / float(delta_t)

    def shard(rank):
        start, stop = sharding.pair_range(rank, len(devs), n_pairs)
        if stop <= start:
            return None
        with torch.cuda.device(devs[rank]):
            B = int(batch_pairs) if batch_pairs else _default_batch(Nx, Ny, itemsize, n_cells * (stop - start), devs[rank], pipelined=False)
            return _sweep_pairs(movie_np, start, stop, cell_alpha, cell_beta, devs[rank], tdtype, smoothing_sigma, guess, scale, rtol,
                                max_it, preconditioner, solver_options, B, direct_mode)

    if len(devs) == 1:
        parts = [shard(0)]
    else:
        parts = sharding.run_on_threads(shard, len(devs))
    parts = [p for p in parts if p is not None]
    moments = sum(p[0] for p in parts)
    functional = sum(p[1] for p in parts)
    converged_last = parts[-1][2]                                        # the shard that holds the last pair (:1202)
    count = float(n_pairs * Nx * Ny)
    shape = (len(alphas), len(betas))
    speed_mean = moments[:, 0] / count
    rem_mean = moments[:, 2] / count
    result_dict = {}
    result_dict['speed_alpha_values'] = speed_alpha_values
    result_dict['remodelling_alpha_values'] = remodelling_alpha_values
    result_dict['speed_means'] = speed_mean.reshape(shape)
    result_dict['speed_variances'] = (moments[:, 1] / count - speed_mean ** 2).reshape(shape)
    result_dict['remodelling_means'] = rem_mean.reshape(shape)
    result_dict['remodelling_variances'] = (moments[:, 3] / count - rem_mean ** 2).reshape(shape)
    result_dict['converged'] = converged_last.reshape(shape)
    result_dict['functional'] = functional.reshape(shape)
    if filename is not None:
        np.save(filename, result_dict)
    return result_dict


# ------------------------------------------------------------------------------------------------
# pre-processing upstream of the flow call (SURVEY.md 8f NEXT-3)
# ------------------------------------------------------------------------------------------------
def _to_device_float(movie, dev, dtype=None):
    """Movie -> contiguous device tensor.  Integer movies become float64 (numpy's true division promotes them the
    same way); float32 / float64 movies keep their type unless ``dtype`` is given."""
    movie = np.ascontiguousarray(np.asarray(movie))
    if dtype is None:
        dtype = movie.dtype if movie.dtype in (np.float32, np.float64) else np.float64
    tdtype, _ = engine.resolve_dtype(dtype)
    return hostio.as_native_tensor(movie).to(dev).to(tdtype)


@_on_device
def apply_adaptive_threshold(movie, window_size=51, threshold=0.0, *, device=None):
    """OpenCV mean adaptive threshold of every frame (optical_flow.py:308-338), on the GPU.

    The movie is mapped to uint8 by ``movie/np.max(movie)*255`` (:328); a pixel is True where it exceeds the
    ``window_size`` x ``window_size`` box mean (replicated border) minus ``ceil(threshold)``.  Returns a bool
    array of the movie's shape.  ``window_size`` must be odd, 3..255 (cv2 asserts odd and > 1)."""
    dev = _device(device)
    frames = _to_device_float(movie, dev)
    return engine.adaptive_threshold(frames, window_size, threshold).cpu().numpy()


@_on_device
def apply_clahe(movie, clipLimit=50000, tile_number=10, *, device=None, verbose=False):
    """OpenCV CLAHE of every frame on ``movie.astype(uint16)`` (optical_flow.py:340-374), on the GPU.

    ``tile_number`` tiles split the columns and ``round(tile_number * Ny/Nx)`` tiles the rows - that is what the
    reference's ``tileGridSize=(tile_number, round(tile_number*aspect_ratio))`` means to OpenCV (:364-366).
    Returns float64 like the reference (which also prints the maximum, :372: here only with ``verbose``)."""
    dev = _device(device)
    movie = np.asarray(movie)
    frames = _to_device_float(movie, dev)
    aspect_ratio = movie.shape[2] / movie.shape[1]
    out = engine.clahe(frames, clipLimit, int(tile_number), int(round(tile_number * aspect_ratio)))
    out = out.to(torch.float64).cpu().numpy()
    if verbose:
        print(np.max(out))
    return out


@_on_device
def subtract_background(movie, background, *, thresholding_sigma=10, dtype='float64', device=None):
    """Background subtraction of ``conduct_optical_flow`` (optical_flow.py:195-198; inline there, a function here):
    ``movie - background`` where the ``sigma=10`` blur of the movie exceeds ``background``, 0 elsewhere."""
    dev = _device(device)
    frames = _to_device_float(movie, dev, dtype)
    blurred = engine.blur(frames, thresholding_sigma)
    return engine.subtract_background(frames, blurred, background).to(torch.float64).cpu().numpy()


# ------------------------------------------------------------------------------------------------
# Liu-Shen Jacobi solver behind the deprecated driver (SURVEY.md 8f NEXT-4)
# ------------------------------------------------------------------------------------------------
def _guess(value, n_pairs, Nx, Ny):
    """Initial guess as a [n_pairs or 1, Nx, Ny] float64 stack: one frame-shaped plane for every pair (the reference,
    :497-499) or one plane per pair (what its driver feeds back when it records iterates, :1458-1460)."""
    plane = np.zeros((Nx, Ny)) if value is None else np.asarray(value, dtype=np.float64)
    if plane.shape == (Nx, Ny):
        return plane[None]
    if plane.shape != (n_pairs, Nx, Ny):
        raise ValueError('initial guess must have the frame shape %r or %r, got %r' % ((Nx, Ny), (n_pairs, Nx, Ny), plane.shape))
    return plane


@_on_device
def liu_shen_optical_flow_jit(movie, delta_x=1.0, delta_t=1.0, alpha=100, remodelling_alpha=1.0, initial_v_x=None,
                              initial_v_y=None, initial_remodelling=None, max_iterations=10, tolerance=1e-9,
                              include_remodelling=True, *, dtype='float64', device=None, batch_pairs=None):
    """Liu-Shen variational flow by ``max_iterations`` Jacobi sweeps per frame pair (optical_flow.py:426-673), on the GPU.

    Same positional arguments and the same 5-tuple ``(v_x, v_y, speed, remodelling, last_iteration_index)`` as the
    reference.  Like there, ``tolerance``, ``remodelling_alpha`` and ``include_remodelling`` have no effect, every
    pair restarts from the initial guess (:497-499) and ``remodelling`` is the initial guess handed back (:499).
    The reference's default guesses are 10x10 zero arrays, which only fit 10x10 frames: ``None`` here means zeros
    of the frame shape.  A guess may also be a ``[T-1, Nx, Ny]`` stack (one plane per pair)."""
    dev = _device(device)
    tdtype, _ = engine.resolve_dtype(dtype)
    movie = np.asarray(movie)
    T, Nx, Ny = movie.shape
    if max_iterations < 1:
        raise ValueError('max_iterations must be at least 1 (the reference returns an unbound iteration index otherwise)')
    n_pairs = T - 1
    guess_x = _guess(initial_v_x, n_pairs, Nx, Ny) * delta_t / delta_x
    guess_y = _guess(initial_v_y, n_pairs, Nx, Ny) * delta_t / delta_x
    all_remodelling = np.broadcast_to(_guess(initial_remodelling, n_pairs, Nx, Ny), (n_pairs, Nx, Ny)).copy()
    out = [np.zeros((n_pairs, Nx, Ny)) for _ in range(3)]
    if n_pairs < 1:
        return out[0], out[1], out[2], all_remodelling, max_iterations - 1
    frames = _to_device_float(movie, dev, 'float32' if tdtype == torch.float32 else 'float64')
    gx = torch.from_numpy(guess_x).to(dev).to(tdtype)
    gy = torch.from_numpy(guess_y).to(dev).to(tdtype)
    itemsize = 4 if tdtype == torch.float32 else 8
    B = int(batch_pairs) if batch_pairs else int(max(1, min(n_pairs, (2 << 30) // (6 * Nx * Ny * itemsize))))
    for k0 in range(0, n_pairs, B):
        k1 = min(k0 + B, n_pairs)
        vx = (gx[k0:k1] if gx.shape[0] > 1 else gx.expand(k1 - k0, Nx, Ny)).contiguous().clone()
        vy = (gy[k0:k1] if gy.shape[0] > 1 else gy.expand(k1 - k0, Nx, Ny)).contiguous().clone()
        fields = engine.liu_shen(frames[k0:k1], frames[k0 + 1:k1 + 1], vx, vy, alpha, max_iterations, delta_x / delta_t)
        for dst, src in zip(out, fields):
            dst[k0:k1] = src.to(torch.float64).cpu().numpy()
    return out[0], out[1], out[2], all_remodelling, max_iterations - 1


def conduct_variational_optical_flow_deprecated(movie, delta_x=1.0, delta_t=1.0, speed_alpha=1.0, remodelling_alpha=1000.0,
                                                v_x_guess=0.1, v_y_guess=0.1, remodelling_guess=0.5, max_iterations=10,
                                                smoothing_sigma=None, return_iterations=False, iteration_stepsize=1,
                                                tolerance=1e-10, include_remodelling=True, use_liu_shen=False, *,
                                                dtype='float64', device=None):
    """Driver of the Liu-Shen solver (optical_flow.py:1318-1529): optional blur, constant initial guesses, optionally
    the iterates every ``iteration_stepsize`` sweeps.

    The reference driver cannot run as committed - it unpacks four values from a function that returns five
    (:1422, :1492 vs :673); this one takes the five and otherwise keeps the reference's result keys, including
    ``total_iterations = max_iterations``."""
    if smoothing_sigma is not None:
        movie_to_analyse = blur_movie(movie, smoothing_sigma, device=device)
    else:
        movie_to_analyse = movie
    if not use_liu_shen:
        raise ValueError("I can currently only really do this for the liu shen jitted method")
    movie_np = np.asarray(movie)
    shape = (movie_np.shape[1], movie_np.shape[2])
    initial_v_x = np.full(shape, float(v_x_guess))
    initial_v_y = np.full(shape, float(v_y_guess))
    initial_remodelling = np.full(shape, float(remodelling_guess))
    solver = dict(tolerance=tolerance, include_remodelling=include_remodelling, dtype=dtype, device=device)
    result = dict()
    if return_iterations:
        recorded = max_iterations // iteration_stepsize
        fields = liu_shen_optical_flow_jit(movie_to_analyse, delta_x, delta_t, speed_alpha, remodelling_alpha, initial_v_x,
                                           initial_v_y, initial_remodelling, max_iterations=iteration_stepsize, **solver)[:4]
        steps = [np.zeros((fields[0].shape[0], recorded + 1) + shape) for _ in range(4)]
        firsts = (initial_v_x, initial_v_y, np.sqrt(initial_v_x ** 2 + initial_v_y ** 2), initial_remodelling)
        for store, first, field in zip(steps, firsts, fields):
            store[:, 0] = first
            store[:, 1] = field
        for index in range(2, recorded + 1):
            # the reference feeds the velocity stacks back as the next initial guess (:1458-1460): one plane per pair
            fields = liu_shen_optical_flow_jit(movie_to_analyse, delta_x, delta_t, speed_alpha, remodelling_alpha, fields[0],
                                               fields[1], fields[3], max_iterations=iteration_stepsize, **solver)[:4]
            for store, field in zip(steps, fields):
                store[:, index] = field
        for key, store in zip(('v_x', 'v_y', 'speed', 'remodelling'), steps):
            result[key] = store[:, -1]
            result[key + '_steps'] = store
        result['iteration_stepsize'] = iteration_stepsize
    else:
        fields = liu_shen_optical_flow_jit(movie_to_analyse, delta_x, delta_t, speed_alpha, remodelling_alpha, initial_v_x,
                                           initial_v_y, initial_remodelling, max_iterations, **solver)
        for key, field in zip(('v_x', 'v_y', 'speed', 'remodelling'), fields):
            result[key] = field
    result['original_data'] = movie
    result['delta_x'] = delta_x
    result['delta_t'] = delta_t
    result['blurred_data'] = movie_to_analyse
    result['max_iterations'] = max_iterations
    result['total_iterations'] = max_iterations
    return result
