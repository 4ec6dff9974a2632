"""Slab decomposition: host-side bookkeeping on the CPU, the solve itself on >= 2 GPUs."""
import ctypes
import json
import os
import subprocess
import sys

import numpy as np
import pytest

from opticalflow_b200 import _lib, slab

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize('Nx,world', [(8192, 8), (8192, 2), (1024, 4), (130, 2), (301, 3), (66, 2), (2048, 8)])
def test_partition_covers_every_row_once(Nx, world):
    R, Lg = slab.partition(Nx, Nx, world)
    assert Lg >= 1 and R % (1 << Lg) == 0
    covered = np.zeros(Nx, dtype=int)
    for rank in range(world):
        first, owned = slab.owned_rows(Nx, R, rank)
        assert owned >= 1 << Lg or rank == world - 1
        assert first % 2 == 1                       # slabs start on odd rows: coarse node K sits on fine row 2K-1
        lo, hi = slab.kept_rows(Nx, R, rank, world)
        covered[lo:hi] += 1
        s = _lib.Slab(rank=rank, world=world, Nx_global=Nx, rows_per_rank=R, gather_level=Lg, reserved=0, comm=None)
        f, o = ctypes.c_int(), ctypes.c_int()
        _lib.check(_lib.load().ofb_slab_local_rows(ctypes.byref(s), ctypes.byref(f), ctypes.byref(o)))
        assert (f.value, o.value) == (first, owned)
    assert (covered == 1).all()


def test_partition_d5_shape():
    """Config d-5: 8192 rows over 8 GPUs = 1024 rows per slab (the last slab owns 1022)."""
    R, Lg = slab.partition(8192, 8192, 8)
    assert R == 1024
    assert slab.owned_rows(8192, R, 7) == (7169, 1022)
    R4, Lg4 = slab.partition(8192, 8192, 8, gather_level=4)
    assert (R4, Lg4) == (1024, 4)


def test_partition_rejects_impossible_splits():
    with pytest.raises(_lib.OfbError):
        slab.partition(12, 12, 8)
    with pytest.raises(_lib.OfbError):
        slab.partition(512, 512, 1)


@pytest.mark.gpu
@pytest.mark.parametrize('cases', ['small', 'medium'])
def test_slab_solve_matches_single_gpu(cases):
    """Two ranks, one GPU each, against the single-GPU solve of the same pair (fp64 1e-6 / fp32 1e-3 rel-L2)."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip('needs two GPUs (run with gpurun --gpus 2)')
    cmd = [sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node', '2', '--master-addr', '127.0.0.1',
           '--master-port', '29517', os.path.join(ROOT, 'tools', 'slab_check.py'), '--cases', cases]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    lines = [json.loads(l) for l in out.stdout.splitlines() if l.startswith('{')]
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    assert lines and all(l['ok'] for l in lines)
    for l in lines:
        assert abs(l['iterations'] - l['iterations_single']) <= 1


@pytest.mark.gpu
def test_slab_solve_matches_the_oracle(tmp_path):
    """The slab-decomposed solve on two ranks against the ORACLE's direct solve of the same pair (the reference's matrix and SuperLU,
    oracle/flow_oracle.py) - not only against the single-GPU solve: a 160 x 1024 pair, float64, rtol 1e-11."""
    import torch
    from oracle import flow_oracle as fo
    if torch.cuda.device_count() < 2:
        pytest.skip('needs two GPUs (run with gpurun --gpus 2)')
    dump = str(tmp_path / 'slab_result.npz')
    cmd = [sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node', '2', '--master-addr', '127.0.0.1',
           '--master-port', '29518', os.path.join(ROOT, 'tools', 'slab_check.py'), '--cases', 'oracle', '--dump', dump]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    z = np.load(dump)
    want = fo.variational_optical_flow(z['movie'], speed_alpha=float(z['alpha']), remodelling_alpha=float(z['beta']),
                                       smoothing_sigma=float(z['sigma']))
    for key in ('v_x', 'v_y', 'speed', 'remodelling'):
        err = np.linalg.norm(z[key] - want[key]) / np.linalg.norm(want[key])
        assert err < 1e-6, (key, err)
