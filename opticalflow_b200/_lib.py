"""ctypes binding of the C ABI in ``include/ofb200.h`` (``libofb200.so``, built by ``__graft_entry__.build()``).

There is no CPU fallback: if the shared library is missing, importing the symbols raises.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIBRARY_PATH = os.path.join(_HERE, 'libofb200.so')

OFB_F32, OFB_F64 = 0, 1
PRECOND_BJACOBI, PRECOND_MG = 0, 1
FLAG_NO_STREAM = 1
FLAG_NO_FUSED_TAIL = 2
FLAG_LEGACY_COARSE_BOUNDARY = 4
FLAG_SLAB_NCCL_HALO = 8
FLAG_TILE_FASTEST = 16
FLAG_PDL = 32
FLAG_STRIP_ITEMS = 64
FLAG_NO_HALF_STEP_EXIT = 128
FLAG_SERIAL_SETUP = 256
FLAG_FUSED_PROLONG = 512
FLAG_POLL_AHEAD = 1024
FLAG_FP32_COEF = 2048
FLAG_BF16_RESIDUAL = 4096
ABI_VERSION = 6

# every symbol include/ofb200.h declares
EXPORTED_SYMBOLS = ('ofb_version', 'ofb_last_error', 'ofb_default_options', 'ofb_launch_count', 'ofb_last_unconverged', 'ofb_last_restarts', 'ofb_workspace_bytes', 'ofb_blur',
                    'ofb_build_rhs', 'ofb_apply', 'ofb_apply_eliminated', 'ofb_smooth', 'ofb_solve', 'ofb_precondition',
                    'ofb_epilogue', 'ofb_slab_partition', 'ofb_slab_local_rows', 'ofb_comm_unique_id', 'ofb_comm_create',
                    'ofb_comm_destroy', 'ofb_comm_peer_halo', 'ofb_slab_workspace_bytes', 'ofb_solve_slab', 'ofb_epilogue_slab',
                    'ofb_liu_shen', 'ofb_movie_max', 'ofb_adaptive_threshold', 'ofb_clahe_workspace_bytes', 'ofb_clahe',
                    'ofb_subtract_background', 'ofb_widen', 'ofb_blur_noalloc', 'ofb_direct_workspace_bytes', 'ofb_direct_solve', 'ofb_direct_graph_state')
BLUR_NOALLOC_MAX_BYTES = 16 << 20


class SolveOptions(ctypes.Structure):
    _fields_ = [('precond', ctypes.c_int), ('nu_pre', ctypes.c_int), ('nu_post', ctypes.c_int),
                ('omega', ctypes.c_double), ('check_every', ctypes.c_int), ('rows_per_strip', ctypes.c_int),
                ('coarse_sweeps', ctypes.c_int), ('flags', ctypes.c_int),
                ('omega_pre', ctypes.c_double * 4), ('omega_post', ctypes.c_double * 4),
                ('max_levels', ctypes.c_int), ('reserved', ctypes.c_int)]


class Slab(ctypes.Structure):
    """ofb_slab: one rank's share of a slab-decomposed frame pair (include/ofb200.h)."""
    _fields_ = [('rank', ctypes.c_int), ('world', ctypes.c_int), ('Nx_global', ctypes.c_int),
                ('rows_per_rank', ctypes.c_int), ('gather_level', ctypes.c_int), ('reserved', ctypes.c_int),
                ('comm', ctypes.c_void_p)]


class NdLevel(ctypes.Structure):
    """ofb_nd_level: one level of the nested-dissection plan (include/ofb200.h); pointers are device addresses."""
    _fields_ = [('n_nodes', ctypes.c_int), ('n_own', ctypes.c_int), ('n_border', ctypes.c_int), ('reserved', ctypes.c_int),
                ('own', ctypes.c_void_p), ('border', ctypes.c_void_p), ('parent_slot', ctypes.c_void_p)]


class OfbError(RuntimeError):
    pass


_lib = None


def load():
    """Load libofb200.so (once).  Raises if it has not been built - the product path never falls back."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.isfile(LIBRARY_PATH):
        raise OfbError('CUDA library %s not found: run `python -c "import __graft_entry__ as g; g.build()"` '
                       '(nvcc, sm_100a). There is no CPU fallback.' % LIBRARY_PATH)
    lib = ctypes.CDLL(LIBRARY_PATH)
    vp, dp, ip, i, d, sz = (ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_double,
                            ctypes.c_size_t)
    lib.ofb_version.restype = i
    lib.ofb_version.argtypes = []
    lib.ofb_last_error.restype = ctypes.c_char_p
    lib.ofb_last_error.argtypes = []
    lib.ofb_default_options.restype = None
    lib.ofb_default_options.argtypes = [ctypes.POINTER(SolveOptions)]
    lib.ofb_launch_count.restype = ctypes.c_longlong
    lib.ofb_launch_count.argtypes = []
    lib.ofb_last_unconverged.restype = i
    lib.ofb_last_unconverged.argtypes = []
    lib.ofb_last_restarts.restype = i
    lib.ofb_last_restarts.argtypes = []
    lib.ofb_workspace_bytes.restype = sz
    lib.ofb_workspace_bytes.argtypes = [i, i, i, i, i]
    lib.ofb_blur.restype = i
    lib.ofb_blur.argtypes = [vp, vp, vp, dp, i, i, i, i, i, vp]
    lib.ofb_build_rhs.restype = i
    lib.ofb_build_rhs.argtypes = [vp, vp, vp, i, i, i, i, vp]
    lib.ofb_apply.restype = i
    lib.ofb_apply.argtypes = [vp, vp, vp, dp, dp, i, i, i, i, vp]
    lib.ofb_apply_eliminated.restype = i
    lib.ofb_apply_eliminated.argtypes = [vp, vp, vp, dp, dp, i, i, i, i, vp]
    lib.ofb_smooth.restype = i
    lib.ofb_smooth.argtypes = [vp, vp, vp, vp, dp, dp, ctypes.POINTER(SolveOptions), i, i, i, i, vp]
    lib.ofb_solve.restype = i
    lib.ofb_solve.argtypes = [vp, vp, vp, dp, dp, d, i, ctypes.POINTER(SolveOptions), vp, sz, ip, dp, ip, dp,
                              i, i, i, i, vp]
    lib.ofb_precondition.restype = i
    lib.ofb_precondition.argtypes = [vp, vp, vp, dp, dp, ctypes.POINTER(SolveOptions), vp, sz, i, i, i, i, vp]
    lib.ofb_epilogue.restype = i
    lib.ofb_epilogue.argtypes = [vp, vp, vp, vp, vp, vp, vp, dp, dp, dp, dp, d, i, i, i, i, vp]
    intp, slabp = ctypes.POINTER(ctypes.c_int), ctypes.POINTER(Slab)
    lib.ofb_slab_partition.restype = i
    lib.ofb_slab_partition.argtypes = [i, i, i, i, intp, intp]
    lib.ofb_slab_local_rows.restype = i
    lib.ofb_slab_local_rows.argtypes = [slabp, intp, intp]
    lib.ofb_comm_unique_id.restype = i
    lib.ofb_comm_unique_id.argtypes = [vp]
    lib.ofb_comm_create.restype = i
    lib.ofb_comm_create.argtypes = [vp, i, i, ctypes.POINTER(ctypes.c_void_p)]
    lib.ofb_comm_destroy.restype = i
    lib.ofb_comm_destroy.argtypes = [vp]
    lib.ofb_comm_peer_halo.restype = i
    lib.ofb_comm_peer_halo.argtypes = [vp]
    lib.ofb_slab_workspace_bytes.restype = sz
    lib.ofb_slab_workspace_bytes.argtypes = [i, i, i, slabp]
    lib.ofb_solve_slab.restype = i
    lib.ofb_solve_slab.argtypes = [vp, vp, vp, dp, dp, d, i, ctypes.POINTER(SolveOptions), vp, sz, ip, dp, ip, dp,
                                   i, i, i, i, slabp, vp]
    lib.ofb_epilogue_slab.restype = i
    lib.ofb_epilogue_slab.argtypes = [vp, vp, vp, vp, vp, vp, vp, dp, dp, dp, dp, d, i, i, i, i, slabp, vp]
    ll = ctypes.c_longlong
    lib.ofb_liu_shen.restype = i
    lib.ofb_liu_shen.argtypes = [vp, vp, vp, vp, vp, vp, vp, d, i, d, i, i, i, i, vp]
    lib.ofb_movie_max.restype = i
    lib.ofb_movie_max.argtypes = [vp, ll, i, dp, vp]
    lib.ofb_adaptive_threshold.restype = i
    lib.ofb_adaptive_threshold.argtypes = [vp, dp, vp, vp, i, i, i, i, i, i, vp]
    lib.ofb_clahe_workspace_bytes.restype = sz
    lib.ofb_clahe_workspace_bytes.argtypes = [i, i, i]
    lib.ofb_clahe.restype = i
    lib.ofb_clahe.argtypes = [vp, vp, d, i, i, vp, sz, i, i, i, i, vp]
    lib.ofb_widen.restype = i
    lib.ofb_widen.argtypes = [vp, vp, ll, vp]
    lib.ofb_subtract_background.restype = i
    lib.ofb_subtract_background.argtypes = [vp, vp, vp, d, ll, i, vp]
    lib.ofb_direct_workspace_bytes.restype = sz
    lib.ofb_direct_workspace_bytes.argtypes = [ctypes.POINTER(NdLevel), i, i, i, i]
    lib.ofb_direct_graph_state.restype = i
    lib.ofb_direct_graph_state.argtypes = []
    lib.ofb_direct_solve.restype = i
    lib.ofb_direct_solve.argtypes = [vp, vp, vp, dp, dp, ctypes.POINTER(NdLevel), i, vp, sz, d, dp, ip, i, i, i, i, vp]
    lib.ofb_blur_noalloc.restype = i
    lib.ofb_blur_noalloc.argtypes = [vp, vp, dp, i, i, i, i, i, vp]
    if lib.ofb_version() != ABI_VERSION:
        raise OfbError('libofb200.so ABI version %d != expected %d' % (lib.ofb_version(), ABI_VERSION))
    _lib = lib
    return lib


def check(rc):
    if rc != 0:
        raise OfbError('ofb200 error %d: %s' % (rc, load().ofb_last_error().decode()))


def default_options():
    opt = SolveOptions()
    load().ofb_default_options(ctypes.byref(opt))
    return opt
