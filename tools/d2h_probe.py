"""Host-egress ceiling of one box: pure device -> host copies into page-locked memory, all ranks at once.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/d2h_probe.py

Every rank copies a 1 GiB device buffer into (a) memory from cudaHostAlloc (torch pin_memory) and (b) page-aligned numpy
memory locked with cudaHostRegister (what hostio.ResultArena hands the drop-in), REPS times back to back after a barrier;
rank 0 prints one JSON line with the per-rank and the aggregate GB/s.  This is the ceiling the end-to-end leg of
bench.py can reach at N ranks: if the aggregate stops growing with N the bytes must go, if it grows linearly something in
the drop-in serialises."""
import json
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from opticalflow_b200 import hostio  # noqa: E402

REPS = 8
NBYTES = 1 << 30


def timed_copies(dst, src, stream, barrier):
    for _ in range(2):
        dst.copy_(src, non_blocking=True)
    barrier()
    t0 = time.perf_counter()
    with torch.cuda.stream(stream):
        for _ in range(REPS):
            dst.copy_(src, non_blocking=True)
    stream.synchronize()
    dt = time.perf_counter() - t0
    barrier()
    return REPS * NBYTES / dt / 1e9


def main():
    rank, world, local = int(os.environ.get('RANK', '0')), int(os.environ.get('WORLD_SIZE', '1')), int(os.environ.get('LOCAL_RANK', '0'))
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        dist.init_process_group('nccl', device_id=dev)

    def barrier():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()

    stream = torch.cuda.Stream(dev)
    src = torch.empty(NBYTES, dtype=torch.uint8, device=dev).random_(0, 255)
    results = {}
    pinned = torch.empty(NBYTES, dtype=torch.uint8, pin_memory=True)
    results['d2h_hostalloc'] = timed_copies(pinned, src, stream, barrier)
    results['h2d_hostalloc'] = timed_copies(src, pinned, stream, barrier)
    t0 = time.perf_counter()
    registered = torch.from_numpy(hostio._page_locked_buffer(NBYTES))
    results['register_seconds_per_GiB'] = time.perf_counter() - t0
    results['d2h_registered'] = timed_copies(registered, src, stream, barrier)
    # two copy streams per rank into the two halves: does one engine saturate the link?
    half = NBYTES // 2
    s2 = torch.cuda.Stream(dev)
    barrier()
    t0 = time.perf_counter()
    for _ in range(REPS):
        with torch.cuda.stream(stream):
            pinned[:half].copy_(src[:half], non_blocking=True)
        with torch.cuda.stream(s2):
            pinned[half:].copy_(src[half:], non_blocking=True)
    stream.synchronize(); s2.synchronize()
    results['d2h_two_streams'] = REPS * NBYTES / (time.perf_counter() - t0) / 1e9
    barrier()
    keys = sorted(results)
    t = torch.tensor([results[k] for k in keys], dtype=torch.float64, device=dev)
    if world > 1:
        gathered = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(gathered, t)
    else:
        gathered = [t]
    if rank == 0:
        table = np.stack([g.cpu().numpy() for g in gathered])           # [rank, key]
        line = dict(ranks=world, host_cores=os.cpu_count(), bytes_per_copy=NBYTES, reps=REPS)
        for k, key in enumerate(keys):
            line[key] = dict(per_rank=[round(float(v), 2) for v in table[:, k]],
                             aggregate=round(float(table[:, k].sum()), 2) if not key.startswith('register') else None)
        try:
            line['numa_nodes'] = len([d for d in os.listdir('/sys/devices/system/node') if d.startswith('node')])
        except OSError:
            line['numa_nodes'] = None
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
