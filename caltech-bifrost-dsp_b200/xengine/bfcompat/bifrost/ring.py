"""``bifrost.ring``: Ring + WriteSpan on the ring shim, with span data presented as compat arrays."""
from xengine import ringshim
from .ndarray import _DevArray, _HostArray


def _wrap(buf, space):
    if hasattr(buf, "data_ptr") or space != "system":
        return _DevArray(buf, space)
    return buf.view(_HostArray)


class _SpanView(object):
    def __init__(self, span, space):
        self._span, self._space = span, space
        self.size = span.size

    @property
    def data(self):
        return _wrap(self._span.data, self._space)

    def data_view(self, dtype="u8", shape=None):
        v = _wrap(self._span.data, self._space)
        v = v.view(dtype) if isinstance(v, _DevArray) else ringshim.view_as(self._span.data, dtype).view(_HostArray)
        return v.reshape(shape) if shape is not None else v

    def close(self, *a):
        self._span.close(*a)

    commit = close

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self._span.close()
        return False


class WriteSpan(_SpanView):
    """``WriteSpan(oseq.ring, nbytes, nonblocking=False)`` (corr_block.py:435)."""

    def __init__(self, ring, size, nonblocking=False):
        base = getattr(ring, "_shim", ring)
        super(WriteSpan, self).__init__(ringshim.WriteSpan(base, size, nonblocking=nonblocking), base.space)


class _WriteSeq(object):
    def __init__(self, seq, ring):
        self._seq, self.ring = seq, ring

    def reserve(self, size, nonblocking=False):
        return _SpanView(self._seq.reserve(size, nonblocking), self.ring.space)

    def end(self):
        self._seq.end()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self._seq.end()
        return False


class _Writer(object):
    def __init__(self, w, ring):
        self._w, self._ring = w, ring

    def __enter__(self):
        self._w.__enter__()
        return self

    def __exit__(self, *exc):
        return self._w.__exit__(*exc)

    def begin_sequence(self, *a, **k):
        return _WriteSeq(self._w.begin_sequence(*a, **k), self._ring)


class _ReadSeq(object):
    def __init__(self, seq, ring):
        self._seq, self._ring = seq, ring
        self.header, self.time_tag, self.nringlet = seq.header, seq.time_tag, seq.nringlet

    def read(self, gulp):
        for sp in self._seq.read(gulp):
            yield _SpanView(sp, self._ring.space)


class Ring(object):
    def __init__(self, space="system", name="ring", **kw):
        self._shim = ringshim.Ring(name, space=space, **kw)
        self.space, self.name = space, name

    def resize(self, *a, **k):
        self._shim.resize(*a, **k)

    def begin_writing(self):
        return _Writer(self._shim.begin_writing(), self)

    def read(self, guarantee=True):
        for seq in self._shim.read(guarantee=guarantee):
            yield _ReadSeq(seq, self)
