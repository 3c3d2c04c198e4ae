"""A small in-process stand-in for the part of bifrost's Ring API the X-engine blocks use.

bifrost is an empty submodule in the reference checkout and cannot be installed here, so the blocks in
``xengine.blocks`` are exercised against this shim; they only use the duck-typed calls listed in
SURVEY.md 8(b) (reference call sites: corr_block.py:342-472, corr_acc_block.py:201-336,
corr_subsel_block.py:258-331, beamform_block.py:371-461, beamform_sum_beams_block.py:198-258) and therefore
run unchanged on real ``bifrost.ring.Ring`` objects.

Model: a ring is an ordered list of sequences; a sequence is a JSON header plus a byte stream made of the
spans its writer committed.  Span memory lives in the ring's ``space``:

    'cuda'       torch uint8 tensor on the current CUDA device
    'cuda_host'  pinned host torch tensor (falls back to pageable memory when no CUDA driver is present)
    'system'     numpy-backed host memory

No arithmetic happens here.  Readers block (condition variable) until data, end of sequence or end of
writing arrive, which is what lets every block run ``main()`` on its own thread like the reference does
(lwa352-pipeline.py:296-302).
"""
import threading

import numpy as np

try:  # torch is the device-memory plumbing of this repo; the shim still imports without it for host rings
    import torch
except Exception:  # pragma: no cover
    torch = None


_NP_OF = {
    "u8": np.uint8, "i8": np.int8, "i16": np.int16, "i32": np.int32, "i64": np.int64, "f32": np.float32,
    "f64": np.float64, "cf32": np.complex64, "cf64": np.complex128,
    "ci32": np.dtype([("re", np.int32), ("im", np.int32)]),
}


def np_dtype(dtype):
    """bifrost dtype strings ('i32', 'ci32', 'cf32', ...) or anything numpy understands -> numpy dtype."""
    if isinstance(dtype, str) and dtype in _NP_OF:
        return np.dtype(_NP_OF[dtype])
    return np.dtype(dtype)


_TORCH_OF = None


def _torch_dtype(npdt):
    global _TORCH_OF
    if _TORCH_OF is None:
        _TORCH_OF = {np.dtype(np.uint8): torch.uint8, np.dtype(np.int8): torch.int8, np.dtype(np.int16): torch.int16,
                     np.dtype(np.int32): torch.int32, np.dtype(np.int64): torch.int64,
                     np.dtype(np.float32): torch.float32, np.dtype(np.float64): torch.float64,
                     np.dtype(np.complex64): torch.complex64, np.dtype(np.complex128): torch.complex128}
    return _TORCH_OF[npdt]


def alloc_bytes(nbytes, space):
    """Raw span memory of ``nbytes`` in ``space`` (see module docstring)."""
    nbytes = int(nbytes)
    if space == "cuda":
        if torch is None or not torch.cuda.is_available():
            raise RuntimeError("ring space 'cuda' needs a CUDA device; there is no CPU stand-in for device rings")
        return torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    if space == "cuda_host":
        if torch is not None:
            # torch's caching host allocator: a block is reused once its previous owner has been dropped
            return torch.empty(nbytes, dtype=torch.uint8, pin_memory=torch.cuda.is_available())
        return np.empty(nbytes, dtype=np.uint8)
    if space == "system":
        return np.empty(nbytes, dtype=np.uint8)
    raise ValueError("unknown ring space %r" % (space,))


def view_as(buf, dtype, shape=None):
    """Reinterpret a uint8 buffer (torch tensor or numpy array) as ``dtype`` (and optionally ``shape``)."""
    npdt = np_dtype(dtype)
    if torch is not None and isinstance(buf, torch.Tensor):
        if npdt.fields is not None:          # 'ci32': expose as int32 [..., 2] (re, im)
            out = buf.view(torch.int32).view(-1, 2)
        else:
            out = buf.view(_torch_dtype(npdt))
    elif npdt.fields is not None:
        out = buf.view(np.int32).reshape(-1, 2)
    else:
        out = buf.view(npdt)
    if shape is not None:
        out = out.reshape(*shape) if torch is not None and isinstance(out, torch.Tensor) else out.reshape(shape)
    return out


class _Header(object):
    """``iseq.header.tostring()`` / ``.tobytes()`` as on a bifrost ReadSequence."""

    def __init__(self, text):
        self._raw = text.encode("utf-8") if isinstance(text, str) else bytes(text)

    def tostring(self):
        return self._raw

    tobytes = tostring


class _Sequence(object):
    def __init__(self, ring, time_tag, header, nringlet):
        self.ring = ring
        self.time_tag = time_tag
        self.header = _Header(header)
        self.nringlet = nringlet
        self.chunks = []        # committed spans: (buffer, nbytes)
        self.ended = False


class Ring(object):
    def __init__(self, name="ring", space="system", maxspans=8, recycle=False):
        self.name = name
        self.space = space
        # recycle: drop a span's memory once every guaranteed reader has moved past it, as a real ring reuses its
        # buffer (long runs; allocators can then hand the block out again).  Off by default: tests read rings after
        # the writers have finished, without a guarantee.
        self._recycle = recycle
        self._cv = threading.Condition()
        self._seqs = []
        self._writing_done = False
        self._gulp = 0
        self._maxspans = maxspans      # back-pressure: committed-but-unread spans a writer may run ahead by
        self._committed = 0            # spans committed so far (serial number of the next one)
        self._consumed = {}            # reader id -> serial number of the next span it will take
        self._open_seq = None

    # ------------------------------------------------------------------ writer side
    def resize(self, contiguous_span, total_span=None, nringlet=1):
        self._gulp = max(self._gulp, int(contiguous_span))

    def begin_writing(self):
        return _RingWriter(self)

    def end_writing(self):
        with self._cv:
            self._writing_done = True
            for s in self._seqs:
                s.ended = True
            self._cv.notify_all()

    # ------------------------------------------------------------------ reader side
    def read(self, guarantee=True):
        i = 0
        rid = object()
        if guarantee:                  # guaranteed readers exert back-pressure on the writer
            with self._cv:
                self._consumed[rid] = 0
        try:
            while True:
                with self._cv:
                    while i >= len(self._seqs) and not self._writing_done:
                        self._cv.wait()
                    if i >= len(self._seqs):
                        return
                    seq = self._seqs[i]
                yield ReadSequence(self, seq, rid)
                i += 1
        finally:                       # a reader that stops must not stall the writer
            with self._cv:
                self._consumed.pop(rid, None)
                self._cv.notify_all()

    # ------------------------------------------------------------------ internals
    def _commit(self, seq, buf, nbytes):
        with self._cv:
            while self._consumed and self._committed - min(self._consumed.values()) >= self._maxspans:
                self._cv.wait(0.05)
            seq.chunks.append((buf, nbytes, self._committed))
            self._committed += 1
            self._cv.notify_all()


class _RingWriter(object):
    def __init__(self, ring):
        self.ring = ring

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.ring.end_writing()
        return False

    def begin_sequence(self, time_tag=0, header="{}", nringlet=1, name="", gulp_nframe=None):
        seq = _Sequence(self.ring, time_tag, header, nringlet)
        with self.ring._cv:
            self.ring._seqs.append(seq)
            self.ring._cv.notify_all()
        return WriteSequence(self.ring, seq)


class WriteSequence(object):
    def __init__(self, ring, seq):
        self.ring = ring
        self._seq = seq
        ring._open_seq = seq

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.end()
        return False

    def reserve(self, size, nonblocking=False):
        return WriteSpan(self.ring, size, nonblocking=nonblocking, _seq=self._seq)

    def commit_buffer(self, buf, nbytes=None):
        """Commit an existing buffer (already in the ring's space) as the next span, zero-copy: what a writer
        achieves on a real ring by producing straight into the span it reserved."""
        self.ring._commit(self._seq, buf, len(buf) if nbytes is None else int(nbytes))

    def end(self):
        with self.ring._cv:
            self._seq.ended = True
            self.ring._cv.notify_all()


class WriteSpan(object):
    """``WriteSpan(oseq.ring, nbytes, nonblocking=False)`` (corr_block.py:435, corr_acc_block.py:312) or
    ``with oseq.reserve(nbytes) as ospan`` (corr_subsel_block.py:293).  The span is committed to the ring's
    currently open sequence on ``close()`` / context exit."""

    def __init__(self, ring, size, nonblocking=False, _seq=None):
        self.ring = ring
        self.size = int(size)
        self._seq = _seq if _seq is not None else ring._open_seq
        self.data = alloc_bytes(self.size, ring.space)
        self._open = True

    def data_view(self, dtype=np.uint8, shape=None):
        return view_as(self.data, dtype, shape)

    def commit(self, size=None):
        self.close(size)

    def close(self, size=None):
        if self._open:
            self._open = False
            self.ring._commit(self._seq, self.data, self.size if size is None else int(size))

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()
        return False


class ReadSequence(object):
    def __init__(self, ring, seq, rid=None):
        self.ring = ring
        self._seq = seq
        self._rid = rid
        self.header = seq.header
        self.time_tag = seq.time_tag
        self.nringlet = seq.nringlet

    def read(self, gulp_size):
        """Yield ReadSpans of ``gulp_size`` bytes (the last one may be short, as in bifrost)."""
        gulp_size = int(gulp_size)
        ring, seq = self.ring, self._seq
        ichunk, pend, pend_bytes = 0, [], 0
        while True:
            with ring._cv:
                while ichunk >= len(seq.chunks) and not seq.ended:
                    ring._cv.wait()
                if ichunk >= len(seq.chunks):
                    break
                buf, nbytes, serial = seq.chunks[ichunk]
                ichunk += 1
                if self._rid in ring._consumed:
                    ring._consumed[self._rid] = serial + 1
                    if ring._recycle and ichunk >= 2:       # the span BEFORE the one handed out now is released
                        low = min(ring._consumed.values()) - 1
                        for j in range(ichunk - 1):
                            if seq.chunks[j][0] is not None and seq.chunks[j][2] < low:
                                seq.chunks[j] = (None, 0, seq.chunks[j][2])
                ring._cv.notify_all()
                if buf is None:                             # recycled before this (late) reader got here
                    continue
            if not pend and nbytes == gulp_size:      # the common case: zero-copy hand-over of a whole span
                yield ReadSpan(buf[:nbytes] if len(buf) != nbytes else buf, nbytes)
                continue
            pend.append(buf[:nbytes])
            pend_bytes += nbytes
            while pend_bytes >= gulp_size:
                whole = _concat(pend)
                yield ReadSpan(whole[:gulp_size], gulp_size)
                rest = whole[gulp_size:]
                pend = [rest] if len(rest) else []
                pend_bytes = len(rest)
        if pend_bytes:
            yield ReadSpan(_concat(pend), pend_bytes)


def _concat(parts):
    if len(parts) == 1:
        return parts[0]
    if torch is not None and isinstance(parts[0], torch.Tensor):
        return torch.cat(parts)
    return np.concatenate(parts)


class ReadSpan(object):
    def __init__(self, data, size):
        self.data = data
        self.size = int(size)

    def data_view(self, dtype=np.uint8, shape=None):
        return view_as(self.data, dtype, shape)

    def release(self):
        pass
