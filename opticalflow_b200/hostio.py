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
import queue
import threading

import numpy as np
import torch

_NATIVE = {np.dtype(np.uint8): torch.uint8, np.dtype(np.int8): torch.int8, np.dtype(np.int16): torch.int16,
           np.dtype(np.int32): torch.int32, np.dtype(np.int64): torch.int64, np.dtype(np.float16): torch.float16,
           np.dtype(np.float32): torch.float32, np.dtype(np.float64): torch.float64,
           np.dtype(np.uint16): torch.uint16, np.dtype(np.bool_): torch.bool}

_pinned_cache = {}
_pinned_lock = threading.Lock()


def pinned_buffer(tag, shape, dtype):
    """Process-wide cache of pinned staging tensors (cudaHostAlloc is slow; the buffers are reused across calls).

    Buffers are keyed by `tag`; a larger request replaces the cached one."""
    n = int(np.prod(shape))
    with _pinned_lock:
        buf = _pinned_cache.get((tag, dtype))
        if buf is None or buf.numel() < n:
            buf = torch.empty(n, dtype=dtype, pin_memory=True)
            _pinned_cache[(tag, dtype)] = buf
    return buf[:n].view(shape)


def as_native_tensor(array):
    """numpy array -> CPU tensor sharing memory where torch knows the dtype, else a float64 copy."""
    array = np.ascontiguousarray(array)
    if array.dtype in _NATIVE and array.dtype.isnative:
        try:
            return torch.from_numpy(array)
        except TypeError:
            pass
    return torch.from_numpy(array.astype(np.float64))


_last_upload = {}


def upload_frames(frames_np, device, tdtype, tag='ingest'):
    """Host frames [F,Nx,Ny] (any numeric dtype) -> device tensor of the compute dtype, through pinned staging."""
    src = as_native_tensor(frames_np)
    if src.dtype == torch.uint16:                       # no CUDA kernels for uint16 in torch: widen losslessly on the host
        src = src.to(torch.int32)
    key = (tag, str(device))
    previous = _last_upload.get(key)
    if previous is not None:
        previous.synchronize()                          # the staging buffer is free once the previous copy has completed
    stage = pinned_buffer(key, tuple(src.shape), src.dtype)
    stage.copy_(src)                                    # multi-threaded host copy into pinned memory
    dev = stage.to(device, non_blocking=True)
    done = torch.cuda.Event()
    done.record(torch.cuda.current_stream(device))
    _last_upload[key] = done
    return dev if dev.dtype == tdtype else dev.to(tdtype)


class EgressWorker:
    """Background thread: waits for a device->pinned copy, then widens the planes into the caller's arrays."""

    def __init__(self):
        self.tasks = queue.Queue()
        self.error = None
        self.thread = threading.Thread(target=self._run, daemon=True)
        self.thread.start()

    def _run(self):
        while True:
            task = self.tasks.get()
            if task is None:
                return
            event, copies, done = task
            try:
                if event is not None:
                    event.synchronize()
                for dst_np, src_t in copies:
                    torch.from_numpy(dst_np).copy_(src_t)
            except Exception as exc:                    # surfaced by close()
                self.error = exc
            finally:
                done.set()

    def submit(self, event, copies):
        """copies: list of (destination numpy view, source CPU tensor).  Returns a threading.Event set when done."""
        done = threading.Event()
        self.tasks.put((event, copies, done))
        return done

    def close(self):
        self.tasks.put(None)
        self.thread.join()
        if self.error is not None:
            raise self.error
