"""Movie ingest / result egress around the device path (SURVEY.md section 8f NEXT-2).

The reference hands numpy arrays in and gets float64 numpy arrays back (optical_flow.py:769, :1189-1205).
At >100 frame pairs/s the copies and the float32 -> float64 widening on the host cost more than the solve,
so they are taken off the critical path:

  * frames go host -> pinned staging -> device in their native dtype (uint8/uint16 microscopy frames are
    widened on the device, not on the host);
  * result planes travel device -> pinned staging in the compute dtype on a dedicated copy stream and are
    widened into the caller's float64 arrays by a worker thread (torch's multi-threaded CPU copy), while the
    device already solves the next chunk of frame pairs.

PyTorch is plumbing here (pinned memory, streams, events); no arithmetic of the hot path happens in it.
"""
import ctypes
import os
import queue
import threading
import time

import numpy as np
import torch

_NATIVE = {np.dtype(np.uint8): torch.uint8, np.dtype(np.int8): torch.int8, np.dtype(np.int16): torch.int16,
           np.dtype(np.int32): torch.int32, np.dtype(np.int64): torch.int64, np.dtype(np.float16): torch.float16,
           np.dtype(np.float32): torch.float32, np.dtype(np.float64): torch.float64,
           np.dtype(np.uint16): torch.uint16, np.dtype(np.bool_): torch.bool}

_copy_pool = None
_copy_pool_lock = threading.Lock()


def copy_pool():
    """Process-wide pool of plain host threads for memory-bound copies/widenings.  numpy releases the GIL inside
    copyto, and these threads block on a queue when idle - unlike OpenMP teams started from several Python
    threads, whose spin-waiting starved the thread that launches the kernels."""
    global _copy_pool
    with _copy_pool_lock:
        if _copy_pool is None:
            import concurrent.futures
            import os
            # several ranks on one host (torchrun) share its cores: OFB_COPY_THREADS, else cores / LOCAL_WORLD_SIZE
            ranks_here = max(1, int(os.environ.get('LOCAL_WORLD_SIZE', '1') or 1))
            workers = int(os.environ.get('OFB_COPY_THREADS', '0') or 0) or max(2, ((os.cpu_count() or 8) - 2) // ranks_here)
            _copy_pool = concurrent.futures.ThreadPoolExecutor(max_workers=workers,
                                                               thread_name_prefix='ofb-copy')
    return _copy_pool


class _HostCopyDesc(ctypes.Structure):
    _fields_ = [('src', ctypes.c_void_p), ('dst', ctypes.c_void_p), ('count', ctypes.c_size_t), ('kind', ctypes.c_int),
                ('reserved', ctypes.c_int)]


_KINDS = {('float32', 'float64'): 1, ('uint8', 'float64'): 2, ('uint16', 'float64'): 3, ('uint8', 'float32'): 4,
          ('uint16', 'float32'): 5, ('float64', 'float32'): 6}
_host_lib = None
_host_lib_tried = False


def host_library():
    """libofb200_host.so (csrc/hostcopy.cpp): a native pool of copy threads, one ctypes call per batch of copies.
    Returns None if it has not been built; the numpy thread pool below is then used (host copies only - the
    compute path has no such alternative)."""
    global _host_lib, _host_lib_tried
    with _copy_pool_lock:
        if not _host_lib_tried:
            _host_lib_tried = True
            path = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'libofb200_host.so')
            if os.path.isfile(path) and os.environ.get('OFB_HOST_NATIVE', '1') != '0':
                lib = ctypes.CDLL(path)
                lib.ofb_host_pool_start.restype = ctypes.c_int
                lib.ofb_host_pool_start.argtypes = [ctypes.c_int, ctypes.c_int]
                lib.ofb_host_copy_many.restype = ctypes.c_int
                lib.ofb_host_copy_many.argtypes = [ctypes.POINTER(_HostCopyDesc), ctypes.c_int, ctypes.c_size_t]
                ranks_here = max(1, int(os.environ.get('LOCAL_WORLD_SIZE', '1') or 1))
                # the widening is bandwidth bound with non-temporal stores: 6-8 threads saturate a socket, more only take
                # cores from the kernel-launching thread (16-core host: 6 threads 205 ms per 128 pairs, 14 threads 240 ms)
                workers = int(os.environ.get('OFB_COPY_THREADS', '0') or 0) or min(8, max(2, ((os.cpu_count() or 8) - 2) // ranks_here))
                lib.ofb_host_pool_start(workers, int(os.environ.get('OFB_HOST_NT', '1') != '0'))
                _host_lib = lib
    return _host_lib


def parallel_copy(pairs, piece_bytes=4 << 20):
    """dst[...] = src (cast) for every (dst, src) pair of equally shaped C-contiguous arrays.  Returns when all bytes
    are written.  Native path: one call into the host copy pool; otherwise ~piece_bytes numpy.copyto pieces on a
    Python thread pool."""
    lib = host_library()
    if lib is not None:
        native, rest = [], []
        for dst, src in pairs:
            kind = 0 if dst.dtype == src.dtype else _KINDS.get((src.dtype.name, dst.dtype.name))
            ok = kind is not None and dst.flags.c_contiguous and src.flags.c_contiguous and dst.size == src.size
            (native if ok else rest).append((dst, src, kind))
        if native:
            descs = (_HostCopyDesc * len(native))()
            for k, (dst, src, kind) in enumerate(native):
                descs[k] = _HostCopyDesc(src.ctypes.data, dst.ctypes.data, src.nbytes if kind == 0 else src.size, kind, 0)
            lib.ofb_host_copy_many(descs, len(native), 1 << 20)
        pairs = [(dst, src) for dst, src, _ in rest]
        if not pairs:
            return
    pool = copy_pool()
    futures = []
    for dst, src in pairs:
        d, s_ = dst.reshape(-1), src.reshape(-1)
        step = max(1, piece_bytes // max(1, d.itemsize))
        for a in range(0, d.size, step):
            futures.append(pool.submit(np.copyto, d[a:a + step], s_[a:a + step], 'unsafe'))
    for f in futures:
        f.result()


def speed_from_velocities(v_x, v_y, out, piece_bytes=4 << 20):
    """out = sqrt(v_x^2 + v_y^2) (optical_flow.py:1191) on the host copy pool, in pieces.  ``speed`` is a function of two
    arrays that cross the host link anyway; computing it here instead of shipping it takes 8 of the 42 MB per 1024^2 pair
    off that link, which is what bounds the call once several GPUs share one host (profiles/r02_d2h_ceiling.txt)."""
    pool = copy_pool()
    a, b, o = v_x.reshape(-1), v_y.reshape(-1), out.reshape(-1)
    step = max(1, piece_bytes // o.itemsize)
    futures = [pool.submit(np.hypot, a[k:k + step], b[k:k + step], o[k:k + step]) for k in range(0, o.size, step)]
    for f in futures:
        f.result()


def _helper_threads(divisor):
    """CPU copy threads for a helper thread: a share of the cores, so that ingest, egress and the kernel-launching
    thread do not oversubscribe the host."""
    import os
    return max(2, (os.cpu_count() or 8) // divisor)


class PinnedPool:
    """Process-wide pool of pinned staging tensors.  cudaHostAlloc is slow, so buffers are kept and reused across
    calls; a call checks buffers out and returns them, so concurrent calls (one host thread per device, or two
    calls on the same device) never share a staging buffer."""

    def __init__(self):
        self.free = []
        self.lock = threading.Lock()

    def acquire(self, shape, dtype):
        nbytes = int(np.prod(shape)) * torch.empty((), dtype=dtype).element_size()
        with self.lock:
            best = None
            for k, buf in enumerate(self.free):
                if buf.numel() >= nbytes and (best is None or buf.numel() < self.free[best].numel()):
                    best = k
            raw = self.free.pop(best) if best is not None else None
        if raw is None:
            raw = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
        view = raw[:nbytes].view(dtype).view(shape)
        view._ofb_raw = raw
        return view

    def release(self, view):
        raw = getattr(view, '_ofb_raw', None)
        if raw is not None:
            with self.lock:
                self.free.append(raw)


pinned_pool = PinnedPool()


def _page_locked_buffer(nbytes):
    """uint8 array of ``nbytes`` whose pages are locked with cudaHostRegister (exact size: torch's pinned allocator rounds
    requests up to a power of two, which would double a 1.07 GB result plane).  Unregistered when the array is collected."""
    import weakref
    page = 4096
    size = -(-max(int(nbytes), 1) // page) * page
    raw = np.empty(size + page, dtype=np.uint8)
    offset = (-raw.ctypes.data) % page
    buf = raw[offset:offset + size]
    cudart = torch.cuda.cudart()
    err = cudart.cudaHostRegister(buf.ctypes.data, size, 0)
    if int(err) != 0:
        raise RuntimeError('cudaHostRegister of %d bytes failed: %s' % (size, err))
    weakref.finalize(raw, cudart.cudaHostUnregister, buf.ctypes.data)      # runs before raw's memory is released
    return buf[:nbytes] if nbytes > 0 else buf[:0]


def _host_ram_bytes():
    try:
        return os.sysconf('SC_PAGE_SIZE') * os.sysconf('SC_PHYS_PAGES')
    except (ValueError, OSError):
        return 64 << 30


def _ranks_on_this_host():
    return max(1, int(os.environ.get('LOCAL_WORLD_SIZE', '1') or 1))


def direct_egress_limit():
    """Bytes of result arrays one call may page-lock for direct egress (OFB_DIRECT_EGRESS_BYTES): by default a quarter of
    the host's RAM shared by the ranks on it, at least 16 GiB and at most 96 GiB - the 500-frame 1024^2 movie of
    config d-3 returns 25 GB of float64."""
    env = os.environ.get('OFB_DIRECT_EGRESS_BYTES')
    if env:
        return int(env)
    return int(min(96 << 30, max(16 << 30, _host_ram_bytes() // 4 // _ranks_on_this_host())))


class ResultArena:
    """Recycles the memory of result arrays the caller has dropped.

    The drop-in returns ~50 MB of float64 per 1024^2 frame pair.  Fresh anonymous memory is zero-filled by the kernel
    page by page on first touch, which at several hundred pairs per second costs as much host bandwidth as the
    widening itself.  `alloc` hands out arrays carved from buffers whose previous arrays are all gone (a weakref
    finalizer on the carrier view returns the buffer), so a caller that processes movie after movie - or calls in a
    loop - writes into pages that are already mapped.  Contents are uninitialised, like numpy.empty.  At most
    `max_idle_bytes` are kept while idle - page-locked buffers included, which are unregistered when they are dropped
    (OFB_RESULT_ARENA_BYTES, default min(48 GiB, RAM / 4 / ranks on this host); 0 disables recycling)."""

    def __init__(self, max_idle_bytes=None):
        if max_idle_bytes is None:
            max_idle_bytes = int(os.environ.get('OFB_RESULT_ARENA_BYTES', min(48 << 30, _host_ram_bytes() // 4 // _ranks_on_this_host())))
        self.max_idle_bytes = int(max_idle_bytes)
        self.free = []                       # idle uint8 buffers
        self.free_pinned = []                # idle page-locked uint8 buffers
        self.lock = threading.Lock()

    def _give_back(self, buf, pinned=False):
        with self.lock:
            pool = self.free_pinned if pinned else self.free
            if sum(b.nbytes for b in self.free) + sum(b.nbytes for b in self.free_pinned) + buf.nbytes <= self.max_idle_bytes:
                pool.append(buf)

    def alloc(self, shape, dtype, pinned=False):
        """``pinned``: page-locked memory (cudaHostAlloc through torch) the device can copy results into directly.
        Pinning costs ~0.1-0.3 s per GB once; recycled buffers stay pinned.  Raises if the memory cannot be pinned."""
        import weakref
        dtype = np.dtype(dtype)
        need = int(np.prod(shape)) * dtype.itemsize
        if not pinned and (self.max_idle_bytes <= 0 or need < (1 << 20)):
            return np.empty(shape, dtype=dtype)
        buf = None
        with self.lock:
            pool = self.free_pinned if pinned else self.free
            best = None
            for k, b in enumerate(pool):
                if need <= b.nbytes <= need + (need >> 2) and (best is None or b.nbytes < pool[best].nbytes):
                    best = k
            if best is not None:
                buf = pool.pop(best)
        if buf is None:
            buf = _page_locked_buffer(need) if pinned else np.empty(need, dtype=np.uint8)
        # numpy collapses .base chains of ndarray views onto the owning array, so the owner the caller's arrays keep
        # alive must be a per-allocation object of our own: a ctypes window onto the buffer.  When the last view of
        # it is gone the finalizer returns the buffer (which the finalizer's argument list keeps alive until then).
        carrier = (ctypes.c_char * need).from_address(buf.ctypes.data)
        weakref.finalize(carrier, self._give_back, buf, pinned)
        return np.frombuffer(carrier, dtype=dtype).reshape(shape)


result_arena = ResultArena()


def as_native_tensor(array):
    """numpy array -> CPU tensor sharing memory where torch knows the dtype, else a float64 copy."""
    array = np.ascontiguousarray(array)
    if array.dtype in _NATIVE and array.dtype.isnative:
        try:
            return torch.from_numpy(array)
        except TypeError:
            pass
    return torch.from_numpy(array.astype(np.float64))


class FrameUploader:
    """Host frames -> device through two pinned staging buffers owned by the call (native dtype on the wire).

    `prefetch(frames)` runs on a helper thread: host copy into pinned memory, then an asynchronous H2D copy on a
    dedicated upload stream - so the next chunk's frames arrive while the device is still solving the current
    chunk (the calling thread is blocked inside the solver, which releases the GIL).  `take(tdtype)` hands the
    prefetched frames to the compute stream."""

    def __init__(self, max_frames, frame_shape, np_dtype, device):
        probe = as_native_tensor(np.zeros((1,), dtype=np_dtype))
        self.wire_dtype = torch.int32 if probe.dtype == torch.uint16 else probe.dtype   # no CUDA kernels for uint16
        self.stages = [pinned_pool.acquire((max_frames,) + tuple(frame_shape), self.wire_dtype) for _ in range(2)]
        self.events = [None, None]
        self.device = device
        self.stream = torch.cuda.Stream(device)
        self.turn = 0
        self.pending = None
        self.error = None

    def _upload(self, frames_np, slot):
        try:
            torch.cuda.set_device(self.device)          # helper thread: a fresh thread has device 0 current
            src = np.ascontiguousarray(frames_np)
            if self.events[slot] is not None:
                self.events[slot].synchronize()         # the staging buffer is free once its previous copy has completed
            stage = self.stages[slot][:src.shape[0]]
            parallel_copy([(stage.numpy(), src)])       # host copy / widening into pinned memory on the copy pool
            with torch.cuda.stream(self.stream):
                dev = stage.to(self.device, non_blocking=True)
                event = torch.cuda.Event()
                event.record(self.stream)
            self.events[slot] = event
            self.result = (dev, event)
        except Exception as exc:
            self.error = exc

    def prefetch(self, frames_np):
        slot = self.turn
        self.turn ^= 1
        self.pending = threading.Thread(target=self._upload, args=(frames_np, slot), daemon=True)
        self.pending.start()

    def take(self, tdtype):
        self.pending.join()
        self.pending = None
        if self.error is not None:
            raise self.error
        dev, event = self.result
        main = torch.cuda.current_stream(self.device)
        main.wait_event(event)
        dev.record_stream(main)
        return dev if dev.dtype == tdtype else dev.to(tdtype)

    def close(self):
        if self.pending is not None:
            self.pending.join()
        self.stream.synchronize()
        for stage in self.stages:
            pinned_pool.release(stage)


class EgressWorker:
    """Background thread: waits for a device->pinned copy, then widens the planes into the caller's arrays."""

    def __init__(self):
        self.tasks = queue.Queue()
        self.error = None
        self.log = []                                   # (seconds waiting for the device copy, seconds widening) per task
        self.thread = threading.Thread(target=self._run, daemon=True)
        self.thread.start()

    def _run(self):
        while True:
            task = self.tasks.get()
            if task is None:
                return
            event, copies, post, done = task
            try:
                t0 = time.perf_counter()
                if event is not None:
                    event.synchronize()
                t1 = time.perf_counter()
                if copies:
                    parallel_copy([(dst_np, src_t.numpy() if isinstance(src_t, torch.Tensor) else src_t) for dst_np, src_t in copies])
                if post is not None:
                    post()
                self.log.append((t1 - t0, time.perf_counter() - t1))
            except Exception as exc:                    # surfaced by close()
                self.error = exc
            finally:
                done.set()

    def submit(self, event, copies, post=None):
        """copies: list of (destination numpy view, source CPU tensor); post: a callable run after them (both after
        `event`, a CUDA event, has completed).  Returns a threading.Event set when done."""
        done = threading.Event()
        self.tasks.put((event, copies, post, done))
        return done

    def close(self):
        self.tasks.put(None)
        self.thread.join()
        if self.error is not None:
            raise self.error
