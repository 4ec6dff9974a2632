"""Slab-decomposed solve vs the single-GPU solve of the same pair (run under torchrun, one rank per GPU).

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
      tools/slab_check.py --cases small
  ... --cases d5 --time         # the 8192^2 pair of config d-5: timing only (no single-GPU comparison by default)

Rank 0 prints one JSON line per case and exits non-zero if a case misses its tolerance.
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from opticalflow_b200 import optical_flow, slab, synthetic, engine  # noqa: E402

CASES = {
    # name: (Nx, Ny, dtype, rtol, gather_level hint, sigma, tolerance rel-L2)
    'small': [(130, 96, 'float64', 1e-11, 0, 2.0, 1e-6), (130, 96, 'float64', 1e-11, 2, 2.0, 1e-6),
              (131, 100, 'float64', 1e-11, 0, 1.5, 1e-6), (258, 200, 'float32', 1e-6, 3, 2.5, 1e-3),
              (300, 260, 'float64', 1e-11, 2, None, 1e-6)],
    'medium': [(1024, 1024, 'float32', 1e-6, 0, 2.5, 1e-3), (1024, 1024, 'float64', 1e-10, 0, 2.5, 1e-6),
               (2048, 1024, 'float32', 1e-6, 3, 2.5, 1e-3)],
    'd5': [(8192, 8192, 'float32', 1e-6, 0, 2.5, 1e-3)],
    # wide and short: what tests/test_slab.py also holds to the oracle's SuperLU solve (--dump)
    'oracle': [(160, 1024, 'float64', 1e-11, 0, 2.5, 1e-6)],
}


def rel(a, b):
    return float(np.linalg.norm((a - b).ravel()) / max(np.linalg.norm(b.ravel()), 1e-300))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--cases', default='small')
    ap.add_argument('--time', action='store_true', help='time the slab solve (device events, max over ranks)')
    ap.add_argument('--no-compare', action='store_true')
    ap.add_argument('--single', action='store_true', help='with --time: also time the whole pair on rank 0 alone')
    ap.add_argument('--gather-level', type=int, default=-1)
    ap.add_argument('--scan-gather', default='', help='with --time: comma list of gather levels to time')
    ap.add_argument('--ab', action='store_true', help='with --time: time both halo transports')
    ap.add_argument('--nccl-halo', action='store_true', help='halo rows by ncclSend/ncclRecv instead of peer stores')
    ap.add_argument('--alpha', type=float, default=1e5)
    ap.add_argument('--beta', type=float, default=1e4)
    ap.add_argument('--repeats', type=int, default=3)
    ap.add_argument('--dump', default='', help='rank 0: save the movie and the gathered slab result of the last case to this .npz')
    args = ap.parse_args()
    rank, world = int(os.environ['RANK']), int(os.environ['WORLD_SIZE'])
    torch.cuda.set_device(int(os.environ.get('LOCAL_RANK', rank)))
    dist.init_process_group('cpu:gloo,cuda:nccl', rank=rank, world_size=world)
    dev = torch.device('cuda', torch.cuda.current_device())
    comm = slab.SlabComm()
    failed = 0
    from opticalflow_b200 import _lib
    options = engine.make_options('mg', flags=_lib.FLAG_SLAB_NCCL_HALO) if args.nccl_halo else None
    for (Nx, Ny, dtype, rtol, hint, sigma, tol) in CASES[args.cases]:
        if args.gather_level >= 0:
            hint = args.gather_level
        movie = synthetic.translating_movie(Nx, Ny, 2, seed=3)
        line = {'case': '%dx%d %s' % (Nx, Ny, dtype), 'world': world, 'sigma': sigma}
        res = slab.variational_optical_flow_slab(movie, speed_alpha=args.alpha, remodelling_alpha=args.beta, smoothing_sigma=sigma,
                                                 dtype=dtype, rtol=rtol, comm=comm, gather_level=hint, solver_options=options,
                                                 gather_result=not (args.no_compare or args.cases == 'd5'))
        info = res['solver_info'] if res is not None else None
        if rank == 0:
            line.update(iterations=int(info['iterations'][-1]), relres=float(info['relative_residual'][-1]),
                        converged=bool(info['converged'][-1]), peer_halo=bool(info['peer_halo']) and not args.nccl_halo, rows_per_rank=info['rows_per_rank'], gather_level=info['gather_level'])
        if args.time:
            tdtype, _ = engine.resolve_dtype(dtype)
            scan = [int(v) for v in args.scan_gather.split(',')] if args.scan_gather else [hint]
            modes = [False, True] if args.ab else [args.nccl_halo]
            line['timings'] = []
            for gl in scan:
                sl = slab.make_slab(comm, Nx, Ny, gl)
                frames = slab.local_frames(movie, sl, sigma, tdtype, dev)
                for nccl_mode in modes:
                    opts = engine.make_options('mg', flags=_lib.FLAG_SLAB_NCCL_HALO) if nccl_mode else None
                    times = []
                    for _ in range(args.repeats + 1):
                        dist.barrier()
                        torch.cuda.synchronize()
                        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                        e0.record()
                        x, it, rr, cv = slab.solve_slab(frames[0], frames[1], sl, args.alpha, args.beta, rtol, 1000, opts)
                        e1.record()
                        torch.cuda.synchronize()
                        t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
                        dist.all_reduce(t, op=dist.ReduceOp.MAX)
                        times.append(float(t[0]))
                    if rank == 0:
                        line['timings'].append({'gather_level': sl.gather_level, 'halo': 'nccl' if nccl_mode or not comm.peer_halo else 'peer',
                                                'solve_ms': min(times[1:]), 'all_ms': [round(v, 3) for v in times[1:]],
                                                'iterations': int(it[0])})
                del frames
            if rank == 0 and args.single:
                # the same pair on one GPU (whole frame), same options: what the slabs are measured against
                fr = torch.from_numpy(movie).to(device=dev, dtype=tdtype)
                fr = engine.blur(fr, sigma) if sigma is not None else fr
                P1, C1 = fr[:1].contiguous(), fr[1:].contiguous()
                b1 = engine.build_rhs(P1, C1)
                st = []
                for _ in range(args.repeats + 1):
                    x1 = torch.zeros_like(b1)
                    torch.cuda.synchronize()
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record()
                    r1 = engine.solve(P1, b1, x1, args.alpha, args.beta, rtol=rtol, max_it=1000)
                    e1.record()
                    torch.cuda.synchronize()
                    st.append(e0.elapsed_time(e1))
                line.update(single_gpu_solve_ms=min(st[1:]))
                del fr, P1, C1, b1, x1
            dist.barrier()
        if rank == 0 and not (args.no_compare or args.cases == 'd5'):
            ref = optical_flow.variational_optical_flow(movie, speed_alpha=args.alpha, remodelling_alpha=args.beta,
                                                        smoothing_sigma=sigma, dtype=dtype, rtol=rtol, device='cuda:0')
            errs = {k: rel(res[k], ref[k]) for k in ('v_x', 'v_y', 'remodelling', 'speed')}
            if sigma is not None:
                errs['blurred_data'] = float(np.abs(res['blurred_data'] - ref['blurred_data']).max())
            line.update(errors=errs, iterations_single=int(ref['solver_info']['iterations'][-1]),
                        functionals=[res['L1_functional'], ref['L1_functional'], res['remodelling_functional'],
                                     ref['remodelling_functional']])
            ok = all(v <= tol for k, v in errs.items() if k != 'blurred_data') and line['converged']
            ok = ok and abs(res['L1_functional'] - ref['L1_functional']) <= 1e-3 * abs(ref['L1_functional']) + 1e-9
            line['ok'] = bool(ok)
            failed += 0 if ok else 1
        if rank == 0:
            print(json.dumps(line), flush=True)
            if args.dump and res is not None and 'v_x' in res:
                np.savez(args.dump, movie=movie, sigma=np.array(np.nan if sigma is None else sigma), alpha=np.array(args.alpha),
                         beta=np.array(args.beta), **{k: res[k] for k in ('v_x', 'v_y', 'speed', 'remodelling')})
    dist.barrier()
    comm.close()
    dist.destroy_process_group()
    sys.exit(1 if failed else 0)


if __name__ == '__main__':
    main()
