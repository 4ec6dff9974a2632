"""Frame-pair sharding across GPUs.

Every frame pair is its own linear system (optical_flow.py:791-1187); the only coupling in the reference is the
warm start (:803-806), which this implementation drops.  So the pairs of a movie are split into contiguous
ranges, one per GPU, with a one-frame overlap between neighbouring shards, and there is NO data-path collective:
results are concatenated on the host.  Two drivers share the partition logic:

  * `solve_on_devices` - one process, one host thread per device (the drop-in's `devices=[...]` argument);
  * `solve_distributed` - one process per GPU under torchrun (`torch.distributed`), results gathered to rank 0
    as host arrays (gloo-compatible: the tests run it with world_size 2 on the CPU with a stand-in solver).
"""
import threading

import numpy as np

FIELD_KEYS = ('v_x', 'v_y', 'speed', 'remodelling')
SUM_KEYS = ('L1_functional', 'remodelling_functional', 'speed_functional')
INFO_ARRAY_KEYS = ('iterations', 'relative_residual', 'recurrence_residual', 'converged', 'fallback', 'speed_functional_per_pair')


def pair_range(rank, world, n_pairs):
    """Contiguous, balanced range [start, stop) of frame pairs for `rank`; earlier ranks take the remainder."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError('bad rank/world: %r/%r' % (rank, world))
    base, extra = divmod(int(n_pairs), world)
    start = rank * base + min(rank, extra)
    return start, start + base + (1 if rank < extra else 0)


def shard_frames(movie, rank, world):
    """Frames a rank needs for its pairs: pairs [s, e) use frames [s, e] (one-frame overlap with the next shard)."""
    start, stop = pair_range(rank, world, movie.shape[0] - 1)
    return movie[start:stop + 1] if stop > start else movie[start:start]


def merge_results(parts, movie, delta_x, delta_t):
    """Concatenate per-shard result dicts (in shard order) into one reference-shaped result dict."""
    parts = [p for p in parts if p is not None]
    if not parts:
        raise ValueError('no shard produced a result')
    result = {}
    # keys a caller left out with outputs=(...) are absent from every shard; the result type follows result_dtype
    for key in FIELD_KEYS:
        if all(key in p for p in parts):
            result[key] = np.concatenate([p[key] for p in parts], axis=0)
    wire = next((result[k].dtype for k in FIELD_KEYS if k in result), np.dtype(np.float64))
    first = parts[0]
    if 'original_data' in first:
        result['original_data'] = np.asarray(movie).astype(wire, copy=False)
    result['delta_x'] = delta_x
    result['delta_t'] = delta_t
    if 'blurred_data' in first:
        if first.get('blurred_is_original', first['blurred_data'] is first.get('original_data')):
            result['blurred_data'] = result['original_data'] if 'original_data' in result else np.asarray(movie).astype(wire, copy=False)
        else:   # shards overlap by one frame: drop the duplicate first frame of every shard but the first
            result['blurred_data'] = np.concatenate([first['blurred_data']] + [p['blurred_data'][1:] for p in parts[1:]], axis=0)
    result['converged'] = bool(parts[-1]['converged'])                   # last pair only, optical_flow.py:1202
    for key in SUM_KEYS:
        result[key] = float(sum(p[key] for p in parts))
    info = {}
    infos = [p.get('solver_info', {}) for p in parts]
    for key in INFO_ARRAY_KEYS:
        if all(key in i for i in infos):
            info[key] = np.concatenate([np.atleast_1d(i[key]) for i in infos])
    info['shards'] = [len(np.atleast_1d(i.get('iterations', ()))) for i in infos]
    for key in ('h2d_bytes', 'd2h_bytes'):
        if all(key in i for i in infos):
            info[key] = int(sum(i[key] for i in infos))
    result['solver_info'] = info
    return result


def run_on_threads(fn, world):
    """fn(rank) on one host thread per rank; returns the list of results, re-raises the first error."""
    parts, errors = [None] * world, []

    def work(rank):
        try:
            parts[rank] = fn(rank)
        except Exception as exc:        # re-raised on the calling thread
            errors.append(exc)

    threads = [threading.Thread(target=work, args=(r,)) for r in range(world)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    if errors:
        raise errors[0]
    return parts


def solve_on_devices(movie, devices, solve_fn, delta_x=1.0, delta_t=1.0):
    """One host thread per device; `solve_fn(frames, device)` returns a result dict for its shard."""
    movie = np.asarray(movie)
    world = len(devices)
    parts, errors = [None] * world, []

    def work(rank):
        try:
            frames = shard_frames(movie, rank, world)
            if frames.shape[0] >= 2:
                parts[rank] = solve_fn(frames, devices[rank])
        except Exception as exc:        # re-raised on the calling thread
            errors.append(exc)

    threads = [threading.Thread(target=work, args=(r,)) for r in range(world)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    if errors:
        raise errors[0]
    return merge_results(parts, movie, delta_x, delta_t)


def solve_distributed(movie, solve_fn, delta_x=1.0, delta_t=1.0, group=None, dst=0):
    """One process per GPU (torch.distributed already initialised).  Every rank solves its shard with
    `solve_fn(frames)`; the shard results are gathered as host objects to `dst`, which returns the merged
    result dict (other ranks return None).  No collective touches the solve itself."""
    import torch.distributed as dist
    movie = np.asarray(movie)
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    frames = shard_frames(movie, rank, world)
    part = solve_fn(frames) if frames.shape[0] >= 2 else None
    if part is not None:                 # original_data is rebuilt from the movie on dst: do not ship it
        same = 'blurred_data' in part and part['blurred_data'] is part.get('original_data')
        part = dict(part, blurred_is_original=same)
        if 'original_data' in part:
            part['original_data'] = None
        if same:
            part['blurred_data'] = None
    gathered = [None] * world if rank == dst else None
    dist.gather_object(part, gathered, dst=dst, group=group)
    if rank != dst:
        return None
    return merge_results(gathered, movie, delta_x, delta_t)


def max_over_ranks(value, group=None, device=None):
    """Maximum of a host scalar over all ranks (timings are reported as the max over ranks)."""
    import torch
    import torch.distributed as dist
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return float(t[0])
