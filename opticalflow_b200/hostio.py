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
    """Host frames -> device through one pinned staging buffer owned by the call (native dtype on the wire)."""

    def __init__(self, max_frames, frame_shape, np_dtype, device):
        probe = as_native_tensor(np.zeros((1,), dtype=np_dtype))
        self.wire_dtype = torch.int32 if probe.dtype == torch.uint16 else probe.dtype   # no CUDA kernels for uint16
        self.stage = pinned_pool.acquire((max_frames,) + tuple(frame_shape), self.wire_dtype)
        self.device = device
        self.last = None

    def upload(self, frames_np, tdtype):
        src = as_native_tensor(frames_np)
        if self.last is not None:
            self.last.synchronize()                     # the staging buffer is free once the previous copy has completed
        stage = self.stage[:src.shape[0]]
        stage.copy_(src)                                # (multi-threaded) host copy / widening into pinned memory
        dev = stage.to(self.device, non_blocking=True)
        self.last = torch.cuda.Event()
        self.last.record(torch.cuda.current_stream(self.device))
        return dev if dev.dtype == tdtype else dev.to(tdtype)

    def close(self):
        if self.last is not None:
            self.last.synchronize()
        pinned_pool.release(self.stage)


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
