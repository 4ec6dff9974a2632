"""Slab decomposition of one very large frame pair over the GPUs of a node.

The reference cannot split one pair: its only scaling knob is host RAM (optical_flow.py:761).  Here every rank
(one process per GPU, torch.distributed already initialised) holds a slab of consecutive image rows.  The solve
is ONE linear system: the ranks exchange a boundary row per field before every stencil sweep, sum their partial
dot products and all-gather the small multigrid levels - all inside ``ofb_solve_slab`` (NCCL over NVLink, see
include/ofb200.h).  This module is the host side: partition, per-rank ingest (blur with a halo of the kernel
radius), the C-ABI calls, and the gather of the result fields to rank 0.

Row bookkeeping (global row indices, ghost rows 0 and Nx-1 included):
  rank r owns interior rows [a, a+owned), a = 1 + r*R;  its local planes are rows [a-1, a+owned+1)  (owned+2 rows);
  the image is additionally held one row further out on both sides (rows [a-2, a+owned+2)), see the header.
"""
import ctypes

import numpy as np
import torch

from . import _lib, engine


def partition(Nx, Ny, world, gather_level=0):
    """(rows_per_rank, gather_level) for an Nx x Ny frame split over `world` ranks (ofb_slab_partition)."""
    R, L = ctypes.c_int(), ctypes.c_int()
    _lib.check(_lib.load().ofb_slab_partition(int(Nx), int(Ny), int(world), int(gather_level), ctypes.byref(R), ctypes.byref(L)))
    return R.value, L.value


def owned_rows(Nx, rows_per_rank, rank):
    """(first owned global row, number of owned rows) of `rank`."""
    first = 1 + rows_per_rank * rank
    return first, int(min(max(Nx - 2 - rows_per_rank * rank, 0), rows_per_rank))


def kept_rows(Nx, rows_per_rank, rank, world):
    """Global rows [lo, hi) of the result fields that rank contributes: its owned rows, plus the frame's ghost row
    on the first and the last rank."""
    first, owned = owned_rows(Nx, rows_per_rank, rank)
    return (0 if rank == 0 else first), (Nx if rank == world - 1 else first + owned)


class SlabComm:
    """NCCL communicator of the library, bootstrapped through torch.distributed (any backend)."""

    def __init__(self, group=None):
        import torch.distributed as dist
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        lib = _lib.load()
        ident = (ctypes.c_ubyte * 128)()
        if self.rank == 0:
            _lib.check(lib.ofb_comm_unique_id(ident))
        box = [bytes(ident)]
        dist.broadcast_object_list(box, src=0, group=group)
        ident = (ctypes.c_ubyte * 128).from_buffer_copy(box[0])
        handle = ctypes.c_void_p()
        _lib.check(lib.ofb_comm_create(ident, self.rank, self.world, ctypes.byref(handle)))
        self.handle = handle
        self.peer_halo = bool(lib.ofb_comm_peer_halo(handle))      # halo rows by peer stores (else ncclSend/ncclRecv)

    def close(self):
        if self.handle:
            _lib.load().ofb_comm_destroy(self.handle)
            self.handle = None


def make_slab(comm, Nx, Ny, gather_level=0):
    R, L = partition(Nx, Ny, comm.world, gather_level)
    return _lib.Slab(rank=comm.rank, world=comm.world, Nx_global=int(Nx), rows_per_rank=R, gather_level=L, reserved=0,
                     comm=comm.handle)


def upload_rows(movie, slab, sigma, tdtype, device):
    """Host -> device copy of the rows this rank needs: its rows +-2, plus the blur radius.  Returns (tensor, meta)."""
    T, Nx, Ny = movie.shape
    first, owned = owned_rows(Nx, slab.rows_per_rank, slab.rank)
    radius = engine.gaussian_weights(sigma)[1] if sigma is not None else 0
    lo, hi = first - 2, first + owned + 2                              # rows wanted
    src_lo, src_hi = max(lo - radius, 0), min(hi + radius, Nx)         # rows read (blur support)
    part = torch.from_numpy(np.ascontiguousarray(movie[:, src_lo:src_hi])).to(device=device, dtype=tdtype, non_blocking=True)
    return part, (lo, hi, src_lo, Nx, owned)


def blur_rows(part, meta, sigma, out=None):
    """Blur the uploaded rows and cut out [T, owned+4, Ny]: local row m is global row a-2+m (rows outside the frame
    are zero and never read).  Rows of `part` whose blur support was cut by the slab edge are dropped here."""
    lo, hi, src_lo, Nx, owned = meta
    if sigma is not None:
        part = engine.blur(part, sigma)
    T, _, Ny = part.shape
    out = torch.zeros((T, owned + 4, Ny), dtype=part.dtype, device=part.device) if out is None else out
    g_lo, g_hi = max(lo, 0), min(hi, Nx)
    out[:, g_lo - lo:g_hi - lo] = part[:, g_lo - src_lo:g_hi - src_lo]
    return out


def local_frames(movie, slab, sigma, tdtype, device):
    """This rank's rows of every frame, blurred, as a device tensor [T, owned+4, Ny]."""
    part, meta = upload_rows(np.asarray(movie), slab, sigma, tdtype, device)
    return blur_rows(part, meta, sigma)


def solve_slab(P_ext, C_ext, slab, speed_alpha, remodelling_alpha, rtol=1e-6, max_it=1000, options=None, x0=None):
    """One pair on this rank's slab.  P_ext/C_ext: [owned+4, Ny] device tensors from `local_frames`.
    Returns (x [1,3,owned+2,Ny], iterations, relres, converged) - the scalars are identical on every rank."""
    lib = _lib.load()
    Nx_loc, Ny = P_ext.shape[0] - 2, P_ext.shape[1]
    dev, code = P_ext.device, engine._code(P_ext)
    P, C = P_ext[1:Nx_loc + 1], C_ext[1:Nx_loc + 1]                    # contiguous row ranges: the local planes
    b = engine.build_rhs(P.unsqueeze(0), C.unsqueeze(0))
    x = torch.zeros_like(b) if x0 is None else x0
    a, be = engine._weights(speed_alpha, 1, dev), engine._weights(remodelling_alpha, 1, dev)
    opt = options if options is not None else engine.make_options('mg')
    nbytes = lib.ofb_slab_workspace_bytes(1, Ny, code, ctypes.byref(slab))
    ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    iters = torch.zeros(1, dtype=torch.int32, device=dev)
    relres = torch.zeros(1, dtype=torch.float64, device=dev)
    conv = torch.zeros(1, dtype=torch.int32, device=dev)
    _lib.check(lib.ofb_solve_slab(engine._ptr(P), engine._ptr(b), engine._ptr(x), engine._ptr(a), engine._ptr(be), float(rtol),
                                  int(max_it), ctypes.byref(opt), engine._ptr(ws), nbytes, engine._ptr(iters), engine._ptr(relres),
                                  engine._ptr(conv), None, 1, Nx_loc, Ny, code, ctypes.byref(slab), engine._stream_ptr(dev)))
    return x, iters, relres, conv


def epilogue_slab(P_ext, C_ext, x, slab, speed_alpha, remodelling_alpha, scale):
    lib = _lib.load()
    Nx_loc, Ny = P_ext.shape[0] - 2, P_ext.shape[1]
    dev, code = P_ext.device, engine._code(P_ext)
    P, C = P_ext[1:Nx_loc + 1], C_ext[1:Nx_loc + 1]
    a, be = engine._weights(speed_alpha, 1, dev), engine._weights(remodelling_alpha, 1, dev)
    out = torch.empty((4, Nx_loc, Ny), dtype=P_ext.dtype, device=dev)
    fun = torch.empty((1, 3), dtype=torch.float64, device=dev)
    _lib.check(lib.ofb_epilogue_slab(engine._ptr(P), engine._ptr(C), engine._ptr(x), engine._ptr(out[0]), engine._ptr(out[1]),
                                     engine._ptr(out[2]), engine._ptr(out[3]), engine._ptr(fun), None, engine._ptr(a), engine._ptr(be),
                                     float(scale), 1, Nx_loc, Ny, code, ctypes.byref(slab), engine._stream_ptr(dev)))
    return out, fun


def variational_optical_flow_slab(movie, delta_x=1.0, delta_t=1.0, speed_alpha=1.0, remodelling_alpha=1000.0, smoothing_sigma=None,
                                  dtype='float32', rtol=1e-6, max_it=1000, group=None, comm=None, gather_level=0,
                                  solver_options=None, device=None, gather_result=True):
    """`variational_optical_flow` (optical_flow.py:715-1210) for frames too large - or too urgent - for one GPU.

    Call on every rank with the same arguments (torch.distributed initialised, one GPU per rank).  Rank 0 returns
    the reference's result dictionary (plus 'solver_info'); the other ranks return None.  With
    gather_result=False nothing is gathered and every rank returns its own rows as device tensors ('rows' = (lo, hi));
    gather_result='local' returns them as host float64 arrays."""
    import torch.distributed as dist
    movie = np.asarray(movie)
    T, Nx, Ny = movie.shape
    tdtype, _ = engine.resolve_dtype(dtype)
    device = torch.device('cuda', torch.cuda.current_device()) if device is None else torch.device(device)
    own_comm = comm is None
    comm = SlabComm(group) if own_comm else comm
    try:
        slab = make_slab(comm, Nx, Ny, gather_level)
        first, owned = owned_rows(Nx, slab.rows_per_rank, comm.rank)
        lo, hi = kept_rows(Nx, slab.rows_per_rank, comm.rank, comm.world)
        frames = local_frames(movie, slab, smoothing_sigma, tdtype, device)
        fields = torch.empty((4, T - 1, hi - lo, Ny), dtype=tdtype, device=device)
        fun = torch.zeros(3, dtype=torch.float64, device=device)
        its, res, cvs = [], [], []
        for k in range(T - 1):
            x, iters, relres, conv = solve_slab(frames[k], frames[k + 1], slab, speed_alpha, remodelling_alpha, rtol, max_it,
                                                solver_options)
            out, f = epilogue_slab(frames[k], frames[k + 1], x, slab, speed_alpha, remodelling_alpha, delta_x / delta_t)
            fields[:, k] = out[:, lo - (first - 1):hi - (first - 1)]
            fun += f[0]
            its.append(iters); res.append(relres); cvs.append(conv)
        info = {'iterations': torch.cat(its).cpu().numpy(), 'relative_residual': torch.cat(res).cpu().numpy(),
                'converged': torch.cat(cvs).cpu().numpy().astype(bool), 'rows_per_rank': slab.rows_per_rank,
                'gather_level': slab.gather_level, 'world': comm.world, 'peer_halo': comm.peer_halo}
        dist.all_reduce(fun, group=group)                              # functionals: sum of the ranks' partial sums
        if gather_result == 'local':        # every rank keeps its own rows, as host float64 arrays like the reference's
            host = fields.cpu().numpy().astype(np.float64)
            return {'rows': (lo, hi), 'v_x': host[0], 'v_y': host[1], 'speed': host[2], 'remodelling': host[3],
                    'functionals': fun.cpu().numpy(), 'solver_info': info}
        if not gather_result:
            return {'rows': (lo, hi), 'fields': fields, 'functionals': fun.cpu().numpy(), 'solver_info': info}
        blurred = frames[:, lo - (first - 2):hi - (first - 2)]
        parts = [None] * comm.world if comm.rank == 0 else None
        dist.gather_object((lo, hi, fields.cpu().numpy(), blurred.cpu().numpy() if smoothing_sigma is not None else None), parts,
                           dst=0, group=group)
        if comm.rank != 0:
            return None
        full = np.empty((4, T - 1, Nx, Ny), dtype=np.float64)
        original = movie.astype(np.float64)
        blurred_full = np.empty_like(original) if smoothing_sigma is not None else original
        for plo, phi, pf, pb in parts:
            full[:, :, plo:phi] = pf
            if pb is not None:
                blurred_full[:, plo:phi] = pb
        fn = fun.cpu().numpy()
        return {'v_x': full[0], 'v_y': full[1], 'speed': full[2], 'remodelling': full[3], 'original_data': original,
                'blurred_data': blurred_full, 'delta_x': delta_x, 'delta_t': delta_t, 'converged': bool(info['converged'][-1]),
                'L1_functional': float(fn[0]), 'speed_functional': float(fn[2]), 'remodelling_functional': float(fn[2]),
                'solver_info': info}
    finally:
        if own_comm:
            comm.close()
