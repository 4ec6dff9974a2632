"""CPU tests of the multi-GPU host logic: pair partition, shard merge, the thread driver and the
torch.distributed driver (gloo, world_size 2) with a stand-in solver - no GPU, no collective on the data path."""
import os
import socket
import sys

import numpy as np
import pytest

from conftest import ROOT


def fake_solve(frames, device=None):
    """Stand-in for the GPU solve: per-pair fields that identify the frames they came from."""
    frames = np.asarray(frames, dtype=np.float64)
    n = frames.shape[0] - 1
    diff = frames[1:] - frames[:-1]
    out = dict(v_x=diff.copy(), v_y=2 * diff, speed=np.abs(diff), remodelling=frames[:-1] + frames[1:],
               original_data=frames, delta_x=1.0, delta_t=1.0, blurred_data=frames + 0.5, converged=True,
               L1_functional=float(diff.sum()), remodelling_functional=float(n), speed_functional=float(n))
    out['solver_info'] = dict(iterations=np.arange(n, dtype=np.int32), relative_residual=np.zeros(n),
                              converged=np.ones(n, bool), fallback=np.zeros(n, bool),
                              speed_functional_per_pair=np.zeros(n), h2d_bytes=frames.nbytes, d2h_bytes=4 * diff.nbytes)
    return out


def test_pair_range_partitions_exactly():
    from opticalflow_b200 import sharding
    for n_pairs in (0, 1, 5, 16, 499):
        for world in (1, 2, 3, 8):
            ranges = [sharding.pair_range(r, world, n_pairs) for r in range(world)]
            assert ranges[0][0] == 0 and ranges[-1][1] == n_pairs
            assert all(ranges[k][1] == ranges[k + 1][0] for k in range(world - 1))
            sizes = [b - a for a, b in ranges]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        sharding.pair_range(2, 2, 10)


@pytest.mark.parametrize('n_frames,world', [(9, 2), (6, 4), (3, 8), (500, 8)])
def test_thread_driver_equals_unsharded(n_frames, world):
    from opticalflow_b200 import sharding
    rng = np.random.default_rng(n_frames)
    movie = rng.uniform(0, 255, (n_frames, 6, 7))
    want = fake_solve(movie)
    got = sharding.solve_on_devices(movie, ['dev%d' % k for k in range(world)], fake_solve)
    for key in sharding.FIELD_KEYS + ('blurred_data', 'original_data'):
        np.testing.assert_array_equal(got[key], want[key])
    assert got['L1_functional'] == pytest.approx(want['L1_functional'])
    assert got['remodelling_functional'] == n_frames - 1
    assert sum(got['solver_info']['shards']) == n_frames - 1
    assert got['solver_info']['iterations'].shape == (n_frames - 1,)


def _gloo_worker(rank, world, port, n_frames, queue):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, 'tests'))
    import torch.distributed as dist
    from opticalflow_b200 import sharding
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        movie = np.random.default_rng(7).uniform(0, 255, (n_frames, 5, 6))
        merged = sharding.solve_distributed(movie, fake_solve)
        slowest = sharding.max_over_ranks(10.0 + rank)
        if rank == 0:
            want = fake_solve(movie)
            ok = all(np.array_equal(merged[k], want[k]) for k in sharding.FIELD_KEYS + ('blurred_data', 'original_data'))
            queue.put((ok, slowest, merged['solver_info']['shards']))
        else:
            assert merged is None
    finally:
        dist.destroy_process_group()


def test_distributed_driver_gloo_world_size_2():
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(('127.0.0.1', 0))
        port = s.getsockname()[1]
    ctx = mp.get_context('spawn')
    queue = ctx.Queue()
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, 8, queue)) for r in range(2)]
    for p in procs:
        p.start()
    ok, slowest, shards = queue.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert ok and slowest == 11.0 and shards == [4, 3]
