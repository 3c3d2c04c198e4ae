"""Host-side runtime services the blocks need besides rings: device selection, core affinity, ProcLog,
device / pinned-host arrays and copies between them.

In the reference these come from bifrost (``bifrost.ndarray``, ``bifrost.device``, ``bifrost.affinity``,
``bifrost.proclog`` -- imports at block_base.py:1-9).  bifrost is absent here, so this module provides the same
handful of calls on top of torch (device memory, streams: plumbing only) and the standard library.  Nothing in
this file does X-engine arithmetic, and nothing silently substitutes host memory for device memory.
"""
import os
import threading

import numpy as np

from . import ringshim
from .ringshim import alloc_bytes, np_dtype, view_as  # noqa: F401

try:
    import torch
except Exception:  # pragma: no cover
    torch = None


# ---------------------------------------------------------------------------------------------- rings
def WriteSpan(ring, size, nonblocking=False):
    """``bifrost.ring.WriteSpan(oseq.ring, nbytes, nonblocking=False)`` (corr_block.py:435, corr_acc_block.py:312) for
    whatever kind of ring the block was given: the in-process shim, or any other ring type through the ``bifrost``
    package that is importable in this process (the real one, or xengine/bfcompat) -- the blocks themselves only use
    the public ring API listed in SURVEY 8(b)."""
    if isinstance(ring, ringshim.Ring):
        return ringshim.WriteSpan(ring, size, nonblocking=nonblocking)
    import bifrost.ring as _bfring
    return _bfring.WriteSpan(ring, size, nonblocking=nonblocking)


def _unwrap(x):
    """Arrays of a foreign ring implementation that wrap a torch tensor / numpy array expose it as ``._xeng_raw``."""
    return getattr(x, "_xeng_raw", x)


# ---------------------------------------------------------------------------------------------- devices
def set_device(gpu):
    """bifrost.device.set_device (BFSetGPU at corr_block.py:236,338)."""
    if gpu is None or gpu < 0:
        return
    if torch is None or not torch.cuda.is_available():
        raise RuntimeError("set_device(%d): no CUDA device available" % gpu)
    torch.cuda.set_device(gpu)


def get_device():
    if torch is None or not torch.cuda.is_available():
        return -1
    return torch.cuda.current_device()


def stream_synchronize():
    """bifrost.device.stream_synchronize (corr_acc_block.py:317): wait for the work queued on this thread's
    current stream.  torch's current stream is the legacy default stream unless the caller changed it, which is
    also the stream libxeng handles use by default (include/xeng.h)."""
    if torch is not None and torch.cuda.is_available():
        torch.cuda.current_stream().synchronize()


# ---------------------------------------------------------------------------------------------- affinity
def set_core(core):
    """bifrost.affinity.set_core: pin the calling thread; -1 = leave unbound (block_base.py:42-43)."""
    if core is None or core < 0:
        return
    try:
        os.sched_setaffinity(0, {int(core)})
    except (AttributeError, OSError):
        pass


def get_core():
    try:
        aff = os.sched_getaffinity(0)
        return min(aff) if len(aff) == 1 else -1
    except AttributeError:
        return -1


# ---------------------------------------------------------------------------------------------- ProcLog
class ProcLog(object):
    """bifrost.proclog.ProcLog: ``update(dict)`` replaces the log's contents (block_base.py:374-387 relies on
    exactly that).  bifrost writes them under /dev/shm/bifrost/<pid>/...; here they are kept in memory and can be
    read back with ``ProcLog.snapshot()`` (what the etcd bridge would publish)."""

    _all = {}
    _lock = threading.Lock()

    def __init__(self, name):
        self.name = name
        self.contents = {}
        with ProcLog._lock:
            ProcLog._all[name] = self

    def update(self, contents):
        self.contents = dict(contents)

    @classmethod
    def snapshot(cls):
        with cls._lock:
            return {k: dict(v.contents) for k, v in cls._all.items()}


# ---------------------------------------------------------------------------------------------- arrays
def empty(shape, dtype, space):
    """``BFArray(shape=..., dtype=..., space=...)``: typed view over freshly allocated memory in ``space``
    ('cuda' | 'cuda_host' | 'system').  'cuda' raises when there is no device."""
    if isinstance(shape, (int, np.integer)):
        shape = (int(shape),)
    npdt = np_dtype(dtype)
    n = int(np.prod(shape)) if len(shape) else 1
    raw = alloc_bytes(n * npdt.itemsize, space)
    if npdt.fields is not None:
        return view_as(raw, npdt, tuple(shape) + (2,))
    return view_as(raw, npdt, tuple(shape))


def zeros(shape, dtype, space):
    a = empty(shape, dtype, space)
    if torch is not None and isinstance(a, torch.Tensor):
        a.zero_()
    else:
        a[...] = 0
    return a


def copy_array(dst, src):
    """bifrost.ndarray.copy_array(dst, src): any combination of device / pinned / pageable memory, enqueued on
    the current stream (callers follow it with stream_synchronize() before committing a span)."""
    dst_in, dst, src = dst, _unwrap(dst), _unwrap(src)
    if torch is not None and isinstance(dst, torch.Tensor):
        if not isinstance(src, torch.Tensor):
            src = torch.from_numpy(np.ascontiguousarray(src))
        dst.view(-1).view(torch.uint8).copy_(src.reshape(-1).view(torch.uint8), non_blocking=True)
        return dst_in
    if torch is not None and isinstance(src, torch.Tensor):
        src = src.detach().cpu().numpy()
    np.copyto(dst.reshape(-1).view(np.uint8), np.ascontiguousarray(src).reshape(-1).view(np.uint8))
    return dst_in


# ---------------------------------------------------------------------------------------------- streams / events
class CopyLane(object):
    """A private stream for copies plus an event per copy: ``copy(dst, src)`` enqueues and returns a ticket,
    ``wait(ticket)`` blocks the calling thread until THAT copy has landed (work queued behind it keeps running).
    On a machine without a CUDA device (host-only tests, 'system' rings) copies run synchronously."""

    def __init__(self):
        self.cuda = torch is not None and torch.cuda.is_available()
        self.stream = torch.cuda.Stream() if self.cuda else None

    def copy(self, dst, src):
        if not self.cuda or not (isinstance(dst, torch.Tensor) and isinstance(src, torch.Tensor)):
            copy_array(dst, src)
            if self.cuda:
                torch.cuda.current_stream().synchronize()
            return None
        with torch.cuda.stream(self.stream):
            dst.view(-1).view(torch.uint8).copy_(src.reshape(-1).view(torch.uint8), non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(self.stream)
        return ev

    @staticmethod
    def wait(ticket):
        if ticket is not None:
            ticket.synchronize()
