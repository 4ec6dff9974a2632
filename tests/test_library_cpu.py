"""CPU tests of the C-ABI library: it builds for sm_100a, loads, and exports every symbol the header declares.
No compute call is made here (no GPU)."""
import ctypes
import os
import re

import pytest

from conftest import ROOT


def test_library_exports_every_declared_symbol(built_library):
    from opticalflow_b200 import _lib
    header = open(os.path.join(ROOT, 'include', 'ofb200.h')).read()
    declared = set(re.findall(r'\b(ofb_[a-z_0-9]+)\s*\(', header))
    declared.discard('ofb_solve_options')
    assert declared == set(_lib.EXPORTED_SYMBOLS)
    for name in declared:
        assert getattr(built_library, name) is not None


def test_version_options_and_workspace(built_library):
    from opticalflow_b200 import _lib
    assert built_library.ofb_version() == _lib.ABI_VERSION
    opt = _lib.default_options()
    assert (opt.precond, opt.nu_pre, opt.nu_post, opt.check_every) == (_lib.PRECOND_MG, 1, 2, 1)
    assert list(opt.omega_pre)[:2] == [1.0, 0.0] and list(opt.omega_post)[:3] == [1.3, 0.5, 0.0]
    assert abs(opt.omega - 0.8) < 1e-15
    small = built_library.ofb_workspace_bytes(1, 64, 64, _lib.OFB_F32, _lib.PRECOND_BJACOBI)
    mg32 = built_library.ofb_workspace_bytes(1, 64, 64, _lib.OFB_F32, _lib.PRECOND_MG)
    mg64 = built_library.ofb_workspace_bytes(1, 64, 64, _lib.OFB_F64, _lib.PRECOND_MG)
    assert 9 * 3 * 64 * 64 * 4 <= small < mg32 < mg64
    assert built_library.ofb_workspace_bytes(4, 64, 64, _lib.OFB_F32, _lib.PRECOND_MG) > 3 * mg32


def test_bad_arguments_are_reported_not_crashed(built_library):
    from opticalflow_b200 import _lib
    assert built_library.ofb_workspace_bytes(1, 4, 64, _lib.OFB_F32, _lib.PRECOND_MG) == 0
    assert b'5x5' in built_library.ofb_last_error()
    rc = built_library.ofb_build_rhs(None, None, None, 0, 64, 64, _lib.OFB_F32, None)
    assert rc < 0 and b'batch' in built_library.ofb_last_error()
    rc = built_library.ofb_solve(None, None, None, None, None, 1e-6, 10, None, None, 0, None, None, None, None,
                                 1, 64, 64, 7, None)
    assert rc < 0 and b'dtype' in built_library.ofb_last_error()


def test_product_path_fails_loudly_without_cuda():
    import numpy as np
    import torch
    from opticalflow_b200 import _lib, optical_flow
    if torch.cuda.is_available():
        pytest.skip('a GPU is present')
    with pytest.raises(_lib.OfbError):
        optical_flow.variational_optical_flow(np.zeros((2, 8, 8)))


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, 'opticalflow_b200')
    for fn in os.listdir(pkg):
        if fn.endswith('.py'):
            text = open(os.path.join(pkg, fn)).read()
            assert 'oracle' not in text.replace('the oracle', ''), fn


def test_result_arena_recycles_only_dropped_arrays():
    """hostio.ResultArena: a buffer goes back to the pool when the LAST view of an array carved from it is gone."""
    import gc
    import numpy as np
    from opticalflow_b200 import hostio
    arena = hostio.ResultArena(max_idle_bytes=1 << 30)
    a = arena.alloc((4, 512, 512), np.float64)
    address = a.ctypes.data
    assert a.flags.c_contiguous and a.flags.writeable and a.dtype == np.float64
    view = a[1:3].reshape(-1)[7:]
    del a
    gc.collect()
    assert not arena.free                                   # a view is still alive
    b = arena.alloc((4, 512, 512), np.float64)
    assert b.ctypes.data != address
    del view
    gc.collect()
    assert len(arena.free) == 1
    c = arena.alloc((4, 512, 512), np.float64)
    assert c.ctypes.data == address                         # recycled
    small = arena.alloc((10, 10), np.float64)               # small arrays are plain numpy.empty
    assert small.base is None
    off = hostio.ResultArena(max_idle_bytes=0)
    assert off.alloc((4, 512, 512), np.float64).base is None


def test_chunk_schedule_covers_every_pair_once_and_tapers_the_end():
    from opticalflow_b200.optical_flow import _chunk_schedule
    for n, B, px in [(128, 16, 1 << 20), (499, 16, 1 << 20), (99, 64, 512 * 512), (1, 16, 1 << 20), (16, 16, 1 << 20),
                     (17, 16, 1 << 20), (40, 16, 64 * 64), (3, 2, 25), (5, 1, 25), (100, 7, 1 << 18)]:
        for taper in (True, False):
            sch = _chunk_schedule(n, B, px, taper=taper)
            assert sch[0][0] == 0 and sch[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(sch, sch[1:])) and all(0 < k1 - k0 <= B for k0, k1 in sch)
    assert [k1 - k0 for k0, k1 in _chunk_schedule(128, 64, 1 << 20)] == [64, 48, 16]       # one short final chunk
    assert [k1 - k0 for k0, k1 in _chunk_schedule(499, 64, 1 << 20)][-3:] == [64, 35, 16]
    assert _chunk_schedule(16, 16, 1 << 20) == [(0, 16)]                 # a single chunk is left alone
    assert [k1 - k0 for k0, k1 in _chunk_schedule(128, 16, 1 << 20)] == [16] * 8           # 16 Mpixel chunks are not split
    assert [k1 - k0 for k0, k1 in _chunk_schedule(40, 16, 64 * 64)] == [16, 16, 8]         # small frames neither
