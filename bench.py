"""bench.py - frame-pairs/sec of the variational optical-flow hot path at 1024^2 (BASELINE.json metric).

  python bench.py [--gpus N --steps K --warmup W] [--impl reference]

A "step" is one batch of B frame pairs of the d-3 workload (1024^2 rotating+shearing textured movie, fp32,
speed_alpha 1e5, remodelling_alpha 1e4, sigma 2.5, rtol 1e-6) through blur -> rhs -> multigrid-BiCGStab solve ->
epilogue, frames already resident in HBM.  Frame pairs are independent, so ranks get disjoint pairs, no data-path
collective ("weak" scaling); the timed region is bracketed by barrier + synchronize and the maximum over ranks is
taken.  `e2e` is the same metric through the public drop-in (host numpy in, host numpy out).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

NX = NY = 1024
KW = dict(speed_alpha=1e5, remodelling_alpha=1e4, smoothing_sigma=2.5)
RTOL = 1e-6
E2E_PAIRS = 128                     # frame pairs of the end-to-end movie (per rank)
METRIC = 'frame-pairs/sec at 1024^2'


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=5)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='ofb200')
    ap.add_argument('--batch', type=int, default=64, help='frame pairs per step per GPU (16/32/64: 775/887/943 pairs/s on one B200)')
    ap.add_argument('--dtype', default='float32')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--no-e2e', action='store_true', help='profiling runs only: skip the end-to-end leg')
    ap.add_argument('--workload', default='d-3', choices=['d-3', 'd-5'],
                    help="d-3: the headline (1024^2 pairs sharded over the GPUs); d-5: ONE 8192^2 pair slab-decomposed over the "
                         "GPUs (needs torchrun with >= 2 ranks)")
    ap.add_argument('--frame', type=int, default=8192, help='d-5: frame size')
    return ap.parse_args()


def workload_movie(n_frames, seed):
    from opticalflow_b200 import synthetic
    return synthetic.rotating_shearing_movie(NX, NY, n_frames, seed=seed).astype(np.float32)


# ------------------------------------------------------------------------------------------ reference / CPU arm
CPU_CROP = 192


def _cpu_pair_task(seed):
    """One bounded sample: a CPU_CROP^2 crop of a d-3 frame pair through the oracle (blur, assembly, spsolve,
    boundary overwrite, functionals) - the reference's own runnable branch restated (oracle/flow_oracle.py)."""
    from oracle import flow_oracle
    from opticalflow_b200 import synthetic
    movie = synthetic.rotating_shearing_movie(CPU_CROP, CPU_CROP, 2, seed=seed)
    t0 = time.perf_counter()
    flow_oracle.variational_optical_flow(movie, **KW)
    return time.perf_counter() - t0


def cpu_throughput(workers, repeats=1):
    """frame-pairs/sec at 1024^2 extrapolated linearly in pixel count from CPU_CROP^2 samples (optimistic for the
    CPU: the sparse direct solve is super-linear), `workers` processes solving independent pairs concurrently."""
    import multiprocessing as mp
    ctx = mp.get_context('spawn')
    with ctx.Pool(workers) as pool:
        pool.map(_cpu_pair_task, range(workers))                  # warm-up: imports, page-in
        t0 = time.perf_counter()
        for r in range(repeats):
            pool.map(_cpu_pair_task, range(100 + r * workers, 100 + (r + 1) * workers))
        dt = time.perf_counter() - t0
    pairs_small = workers * repeats
    scale = (CPU_CROP * CPU_CROP) / float(NX * NY)
    return pairs_small * scale / dt, dt


def run_reference(args):
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    workers = max(1, min(os.cpu_count() or 1, 32))
    for _ in range(args.warmup):
        pass                                                       # warm-up happens inside cpu_throughput's first map
    values = []
    t_all = time.perf_counter()
    for _ in range(args.steps):
        v, _ = cpu_throughput(workers, repeats=1)
        values.append(v)
        if time.perf_counter() - t_all > 240:
            break
    value = float(np.mean(values))
    sample = ('%d concurrent %dx%d crops of d-3 frame pairs per step, oracle port (numpy assembly + SuperLU spsolve), '
              'scaled by pixel count to 1024^2' % (workers, CPU_CROP, CPU_CROP))
    line = dict(metric=METRIC, value=value, unit='frame-pairs/s', n_gpus=args.gpus, steps=len(values), warmup=args.warmup,
                ms_per_step=1e3 * workers * (CPU_CROP ** 2 / float(NX * NY)) / value, higher_is_better=True,
                scaling='weak', vs_baseline=None, dtype='f64', data='synthetic', impl='reference',
                config=dict(workload='d-3: 1024^2 rotating+shearing textured movie, alpha 1e5, beta 1e4, sigma 2.5'),
                cpu_baseline=dict(value=value, unit='frame-pairs/s', cores=workers, kind='port', sample=sample),
                e2e=dict(value=value, unit='frame-pairs/s', h2d_bytes_per_step=0, d2h_bytes_per_step=0),
                gpu_launches=0)
    print(json.dumps(line))


# ------------------------------------------------------------------------------------------ GPU arm
class ClockSampler(threading.Thread):
    """Samples nvidia-smi clocks / throttle reasons of one GPU while the timed region runs."""
    Q = ('clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
         'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.stop_flag = index, [], threading.Event()

    def run(self):
        # ONE nvidia-smi process in loop mode (the profiling recipe's clocks line): spawning one per sample cost the
        # host ~50 ms of CPU each and slowed the end-to-end leg it was sampling
        try:
            proc = subprocess.Popen(['nvidia-smi', '-i', str(self.index), '--query-gpu=' + self.Q, '--format=csv,noheader,nounits',
                                     '-lms', '200'], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            return
        self.stop_flag.wait()
        time.sleep(0.25)                      # let the sample of the last interval arrive
        proc.terminate()
        try:
            out, _ = proc.communicate(timeout=5)
        except Exception:
            proc.kill()
            out = ''
        for line in out.splitlines():
            self.samples.append([f.strip() for f in line.strip().split(',')])

    def summary(self):
        sm = [float(s[0]) for s in self.samples if len(s) >= 7 and s[0].replace('.', '').isdigit()]
        mx = [float(s[1]) for s in self.samples if len(s) >= 7 and s[1].replace('.', '').isdigit()]
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        reasons = sorted({names[k] for s in self.samples if len(s) >= 7 for k in range(4) if s[3 + k] == 'Active'})
        return dict(sm_mhz=float(np.median(sm)) if sm else None, sm_max_mhz=max(mx) if mx else None, reasons=reasons,
                    samples=len(sm))


def run_gpu(args):
    import torch
    import torch.distributed as dist
    from opticalflow_b200 import _lib, engine, optical_flow

    rank = int(os.environ.get('RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:       # one process per GPU shares the host cores: do not oversubscribe them with CPU copy threads
        torch.set_num_threads(max(2, (os.cpu_count() or 8) // world))
    if world > 1:
        dist.init_process_group('nccl', device_id=dev)
    lib = _lib.load()
    B = args.batch
    tdtype, _ = engine.resolve_dtype(args.dtype)
    itemsize = 4 if args.dtype == 'float32' else 8

    movie = workload_movie(B + 1, seed=2 + rank)                   # each rank: its own disjoint pairs
    frames_dev = torch.from_numpy(movie).to(dev).to(tdtype)
    workspace = engine.Workspace()
    opts = engine.make_options('mg')
    alpha = torch.full((B,), KW['speed_alpha'], dtype=torch.float64, device=dev)
    beta = torch.full((B,), KW['remodelling_alpha'], dtype=torch.float64, device=dev)
    blurred = torch.empty_like(frames_dev)
    b = torch.empty((B, 3, NX, NY), dtype=tdtype, device=dev)
    x = torch.empty_like(b)
    fields = torch.empty((4, B, NX, NY), dtype=tdtype, device=dev)
    stats = {}

    def device_step():
        engine.blur(frames_dev, KW['smoothing_sigma'], out=blurred)
        P, C = blurred[:-1], blurred[1:]
        engine.build_rhs(P, C, out=b)
        x.zero_()
        it, rr, cv = engine.solve(P, b, x, alpha, beta, RTOL, 1000, opts, workspace)
        engine.epilogue(P, C, x, alpha, beta, 1.0, out=fields)
        stats['iters'], stats['conv'] = it, cv

    def barrier():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    for _ in range(max(args.warmup, 3)):
        device_step()
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    launches0 = lib.ofb_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        device_step()
    e1.record()
    barrier()
    launches = lib.ofb_launch_count() - launches0
    ms = e0.elapsed_time(e1)
    iters = stats['iters'].cpu().numpy()
    assert int(stats['conv'].sum()) == B, 'not every pair converged: %s' % stats['conv'].tolist()

    # ---- e2e: the public drop-in, host numpy in / host float64 numpy out, every copy inside the timed region.
    # A 129-frame movie (the B+1 generated frames played forwards and backwards: every pair is a valid d-3 pair) so
    # that the drop-in's chunk pipeline (solve chunk k+1 while chunk k's planes go device -> host and are widened) is
    # exercised the way a 500-frame movie would exercise it; 128 pairs keep the float64 results of two calls (the
    # previous one is alive while the next is computed) at 12 GB per rank.
    zigzag = list(range(0, B + 1))
    while len(zigzag) < E2E_PAIRS + 1:
        zigzag += list(range(B - 1, -1, -1)) + list(range(1, B + 1))
    movie_e2e = np.ascontiguousarray(movie[np.array(zigzag[:E2E_PAIRS + 1])])
    pairs_e2e = movie_e2e.shape[0] - 1

    def e2e_step():
        return optical_flow.variational_optical_flow(movie_e2e, dtype=args.dtype, rtol=RTOL, device=dev, **KW)      # the drop-in's own chunking
    if args.no_e2e:
        sampler.stop_flag.set()
        sampler.join()
        e2e_s, h2d, d2h = float('nan'), 0, 0
    else:
        for _ in range(2):
            res = e2e_step()
        barrier()
        t0 = time.perf_counter()
        n_e2e = max(2, min(args.steps, 3))
        for _ in range(n_e2e):
            res = e2e_step()
        barrier()
        e2e_s = (time.perf_counter() - t0) / n_e2e
        sampler.stop_flag.set()
        sampler.join()
        assert res['solver_info']['converged'].all()
        h2d = res['solver_info']['h2d_bytes'] * B // pairs_e2e          # per step of B pairs
        d2h = res['solver_info']['d2h_bytes'] * B // pairs_e2e

    # ---- roofline of the dominant kernel: the level-0 block-Jacobi smoother sweep (10 words / pixel)
    xin = torch.randn_like(b)
    P = blurred[:-1]
    for _ in range(3):
        engine.smooth(P, b, xin, alpha, beta, out=x)
    torch.cuda.synchronize(dev)
    s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 20
    s0.record()
    for _ in range(reps):
        engine.smooth(P, b, xin, alpha, beta, out=x)
    s1.record()
    torch.cuda.synchronize(dev)
    sweep_ms = s0.elapsed_time(s1) / reps
    alg_bytes = 10 * itemsize * NX * NY * B
    achieved = alg_bytes / (sweep_ms * 1e-3) / 1e9
    peaks_file = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.isfile(peaks_file):
        peak, peak_src = float(json.load(open(peaks_file))['hbm_gbs']), 'MEASURED_PEAKS.json hbm_gbs'
    else:
        peak, peak_src = 6650.0, 'fallback (B200_PROFILING.md)'

    traffic = None                                                  # DRAM bytes per launch from the committed ncu --set full capture
    tf = os.path.join(ROOT, 'profiles', 'r01_smoother_traffic.json')
    if os.path.isfile(tf) and args.dtype == 'float32':
        captured = json.load(open(tf)).get(str(B))               # keyed by pairs per launch
        traffic = int(captured['traffic_bytes_per_launch']) if captured else None

    # ---- reduce over ranks
    t = torch.tensor([ms, e2e_s * 1e3], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_max, e2e_ms_max = float(t[0]), float(t[1])
    pairs_total = B * world
    value = pairs_total * args.steps / (ms_max * 1e-3)
    e2e_value = pairs_e2e * world / (e2e_ms_max * 1e-3)

    if rank == 0:
        line = dict(metric=METRIC, value=value, unit='frame-pairs/s', n_gpus=world, steps=args.steps, warmup=max(args.warmup, 3),
                    ms_per_step=ms_max / args.steps, higher_is_better=True, scaling='weak', vs_baseline=None,
                    dtype='f32' if args.dtype == 'float32' else 'f64', data='synthetic',
                    config=dict(workload='d-3: 1024^2 rotating+shearing textured movie, alpha 1e5, beta 1e4, sigma 2.5, '
                                         'rtol 1e-6, BiCGStab + V(1,2) multigrid (block-Jacobi smoother)',
                                pairs_per_step_per_gpu=B, parallelism='frame pairs sharded over %d GPU(s), no collective' % world,
                                l2='working set %.1f GB per step exceeds the 126 MB L2' % (12 * 3 * NX * NY * itemsize * B / 1e9),
                                iterations_per_pair=float(iters.mean()), iterations_max=int(iters.max())),
                    e2e=dict(value=e2e_value, unit='frame-pairs/s', h2d_bytes_per_step=int(h2d), d2h_bytes_per_step=int(d2h)),
                    gpu_launches=int(launches),
                    roofline=dict(bound='hbm', kernel='k_stream<float,true,OpSmooth> level 0 (block-Jacobi smoother, streamed sweep)', achieved=achieved,
                                  peak=peak, unit='GB/s', frac=achieved / peak, traffic=traffic, peak_source=peak_src,
                                  ms_per_launch=sweep_ms, algorithmic_bytes_per_launch=alg_bytes),
                    clocks=sampler.summary())
        if not args.no_cpu_baseline and world == 1:
            workers = max(1, min(os.cpu_count() or 1, 32))
            v, dt = cpu_throughput(workers)
            line['cpu_baseline'] = dict(value=v, unit='frame-pairs/s', cores=workers, kind='port',
                                        sample='%d concurrent %dx%d crops of d-3 pairs (%.1f s), oracle port, scaled by pixel '
                                               'count to 1024^2' % (workers, CPU_CROP, CPU_CROP, dt))
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def run_slab(args):
    """Config d-5: one 8192^2 frame pair, slab-decomposed over the ranks (halo rows by peer stores, dot products
    all-reduced, coarse multigrid levels replicated).  A step = blur -> rhs -> solve -> epilogue of that pair."""
    import torch
    import torch.distributed as dist
    from opticalflow_b200 import _lib, engine, slab, synthetic
    rank, world, local = int(os.environ.get('RANK', '0')), int(os.environ.get('WORLD_SIZE', '1')), int(os.environ.get('LOCAL_RANK', '0'))
    if world < 2:
        raise SystemExit('--workload d-5 needs torchrun with at least 2 ranks (one pair is split over the GPUs)')
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    dist.init_process_group('cpu:gloo,cuda:nccl', rank=rank, world_size=world)
    lib = _lib.load()
    N = args.frame
    tdtype, _ = engine.resolve_dtype(args.dtype)
    itemsize = 4 if args.dtype == 'float32' else 8
    movie = synthetic.translating_movie(N, N, 2, seed=3).astype(np.float32)
    comm = slab.SlabComm()
    sl = slab.make_slab(comm, N, N, 0)
    sigma = KW['smoothing_sigma']
    a = torch.full((1,), KW['speed_alpha'], dtype=torch.float64, device=dev)
    be = torch.full((1,), KW['remodelling_alpha'], dtype=torch.float64, device=dev)
    part, meta = slab.upload_rows(movie, sl, sigma, tdtype, dev)
    stats = {}

    def device_step():
        frames = slab.blur_rows(part, meta, sigma)
        x, it, rr, cv = slab.solve_slab(frames[0], frames[1], sl, a, be, RTOL, 1000)
        out, fun = slab.epilogue_slab(frames[0], frames[1], x, sl, a, be, 1.0)
        stats.update(it=it, cv=cv, frames=frames)

    def barrier():
        torch.cuda.synchronize(dev)
        dist.barrier()
        torch.cuda.synchronize(dev)

    for _ in range(max(args.warmup, 3)):
        device_step()
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    launches0 = lib.ofb_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        device_step()
    e1.record()
    barrier()
    launches = lib.ofb_launch_count() - launches0
    ms = e0.elapsed_time(e1)
    assert int(stats['cv'][0]) == 1, 'the pair did not converge'

    def e2e_step():
        return slab.variational_optical_flow_slab(movie, speed_alpha=KW['speed_alpha'], remodelling_alpha=KW['remodelling_alpha'], smoothing_sigma=sigma, dtype=args.dtype,
                                                  rtol=RTOL, comm=comm, gather_result='local')
    res = e2e_step()
    barrier()
    t0 = time.perf_counter()
    n_e2e = 2
    for _ in range(n_e2e):
        res = e2e_step()
    barrier()
    e2e_s = (time.perf_counter() - t0) / n_e2e
    sampler.stop_flag.set()
    sampler.join()
    lo, hi = res['rows']
    h2d = int(part.numel() * itemsize)
    d2h = int(4 * (hi - lo) * N * itemsize)

    # roofline: the level-0 smoother sweep on this rank's slab
    frames = stats['frames']
    Nx_loc = frames.shape[1] - 2
    P = frames[0][1:Nx_loc + 1].unsqueeze(0)
    bb = torch.randn((1, 3, Nx_loc, N), dtype=tdtype, device=dev)
    xin, xo = torch.randn_like(bb), torch.empty_like(bb)
    for _ in range(3):
        engine.smooth(P, bb, xin, a, be, out=xo)
    torch.cuda.synchronize(dev)
    s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s0.record()
    for _ in range(20):
        engine.smooth(P, bb, xin, a, be, out=xo)
    s1.record()
    torch.cuda.synchronize(dev)
    sweep_ms = s0.elapsed_time(s1) / 20
    alg_bytes = 10 * itemsize * Nx_loc * N
    achieved = alg_bytes / (sweep_ms * 1e-3) / 1e9
    peaks_file = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.isfile(peaks_file):
        peak, peak_src = float(json.load(open(peaks_file))['hbm_gbs']), 'MEASURED_PEAKS.json hbm_gbs'
    else:
        peak, peak_src = 6650.0, 'fallback (B200_PROFILING.md)'
    t = torch.tensor([ms, e2e_s * 1e3], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    tot = torch.tensor([float(h2d), float(d2h)], dtype=torch.float64, device=dev)
    dist.all_reduce(tot)
    if rank == 0:
        line = dict(metric='frame-pairs/sec at %d^2 (one pair, slab-decomposed)' % N, value=args.steps / (float(t[0]) * 1e-3),
                    unit='frame-pairs/s', n_gpus=world, steps=args.steps, warmup=max(args.warmup, 3), ms_per_step=float(t[0]) / args.steps,
                    higher_is_better=True, scaling='strong', vs_baseline=None, dtype='f32' if args.dtype == 'float32' else 'f64',
                    data='synthetic',
                    config=dict(workload='d-5: one %d^2 textured pair shifted by (1.5, -0.7) px, alpha 1e5, beta 1e4, sigma 2.5, rtol 1e-6; '
                                         '%d slabs of %d rows, halo rows by %s, multigrid levels >= %d replicated'
                                         % (N, world, sl.rows_per_rank, 'peer stores (CUDA IPC mailboxes)' if comm.peer_halo else
                                            'ncclSend/ncclRecv', sl.gather_level),
                                parallelism='1-D slab decomposition over %d GPUs' % world, iterations=int(stats['it'][0]),
                                l2='working set per rank %.1f GB exceeds the 126 MB L2' % (12 * 3 * Nx_loc * N * itemsize / 1e9)),
                    e2e=dict(value=1.0 / (float(t[1]) * 1e-3), unit='frame-pairs/s', h2d_bytes_per_step=int(tot[0]),
                             d2h_bytes_per_step=int(tot[1])),
                    gpu_launches=int(launches),
                    roofline=dict(bound='hbm', kernel='k_stream<float,true,OpSmooth> level 0 on one slab', achieved=achieved, peak=peak,
                                  unit='GB/s', frac=achieved / peak, traffic=None, peak_source=peak_src, ms_per_launch=sweep_ms,
                                  algorithmic_bytes_per_launch=alg_bytes),
                    clocks=sampler.summary())
        print(json.dumps(line))
    dist.barrier()
    comm.close()
    dist.destroy_process_group()


if __name__ == '__main__':
    a = parse()
    if a.impl == 'reference':
        run_reference(a)
    elif a.workload == 'd-5':
        run_slab(a)
    else:
        run_gpu(a)
