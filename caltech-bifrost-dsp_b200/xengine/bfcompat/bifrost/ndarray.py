"""``bifrost.ndarray``: typed arrays in 'system' / 'cuda_host' / 'cuda' space with ``as_BFarray()``.
'system' arrays are numpy subclasses (the reference indexes and assigns them like numpy arrays, corr_block.py:259-261,
329); device and pinned arrays wrap the torch tensors of the ring shim."""
import numpy as np

from xengine import ringshim
from xengine import runtime as rt

try:
    import torch
except Exception:  # pragma: no cover
    torch = None


class _Desc(object):
    """What ``as_BFarray()`` hands to a ``_bf.*`` call: an address and a shape."""

    def __init__(self, data, shape):
        self.data, self.shape = data, tuple(shape)


class _HostArray(np.ndarray):
    space = "system"

    def as_BFarray(self):
        return _Desc(self.ctypes.data, self.shape)

    def copy(self, space=None, order="C"):
        if space in (None, "system"):
            return np.array(self, copy=True).view(_HostArray)
        out = ndarray(shape=self.shape, dtype=self.dtype, space=space)
        copy_array(out, self)
        return out


class _DevArray(object):
    """A torch tensor ('cuda' or pinned 'cuda_host'; plain host memory in GPU-less tests) seen as a bifrost array.
    'ci32' is one element to bifrost and a trailing (re, im) axis of int32 here: hidden from shape / reshape."""

    def __init__(self, t, space, bfdtype=None):
        self.t, self.space, self.bfdtype = t, space, bfdtype

    def _hide(self):
        return 1 if self.bfdtype == "ci32" else 0

    @property
    def shape(self):
        s = tuple(self.t.shape)
        return s[:-1] if self._hide() else s

    dtype = property(lambda self: self.t.dtype)
    data = property(lambda self: self.t)        # `arr.data[v] = x` (corr_subsel_block.py:249-250)
    _xeng_raw = property(lambda self: self.t)   # xengine.runtime unwraps foreign arrays through this

    def data_ptr(self):
        return self.t.data_ptr() if hasattr(self.t, "data_ptr") else self.t.ctypes.data

    def as_BFarray(self):
        return _Desc(self.data_ptr(), self.shape)

    def reshape(self, *shape):
        if len(shape) == 1 and not isinstance(shape[0], (int, np.integer)):
            shape = tuple(shape[0])
        shape = tuple(int(x) for x in shape) + ((2,) if self._hide() else ())
        return _DevArray(self.t.reshape(*shape), self.space, self.bfdtype)

    def view(self, dtype):
        flat = self.t.reshape(-1)
        flat = flat.view(torch.uint8) if hasattr(flat, "data_ptr") else flat.view(np.uint8)
        return _DevArray(ringshim.view_as(flat, dtype), self.space, dtype if isinstance(dtype, str) else None)

    def copy(self, space=None, order="C"):
        if space in (None, self.space):
            return _DevArray(self.t.clone(), self.space, self.bfdtype)
        if space == "system":
            h = self.t.detach().cpu().numpy() if hasattr(self.t, "detach") else np.array(self.t)
            return h.view(_HostArray)
        out = ndarray(shape=self.shape, dtype=self.bfdtype or self.t.dtype, space=space)
        copy_array(out, self)
        return out

    def __getitem__(self, k):
        return _DevArray(self.t[k], self.space, self.bfdtype)

    def __setitem__(self, k, v):
        if isinstance(v, _DevArray):
            v = v.t
        if torch is not None and isinstance(self.t, torch.Tensor):
            if not isinstance(v, torch.Tensor):
                v = torch.as_tensor(np.asarray(v))
            v = v.to(self.t.device)
        elif torch is not None and isinstance(v, torch.Tensor):
            v = v.detach().cpu().numpy()
        self.t[k] = v


def ndarray(base=None, space=None, shape=None, dtype=None, **kw):
    """``BFArray(base, dtype=, space=)`` / ``BFArray(shape=, dtype=, space=)`` (corr_block.py:251-261,
    corr_subsel_block.py:210-215, beamform_block.py:238-241)."""
    space = space or "system"
    if base is not None and shape is None:
        b = np.asarray(base, dtype=None if dtype is None else ringshim.np_dtype(dtype))
        if space == "system":
            return np.ascontiguousarray(b).view(_HostArray)
        out = ndarray(shape=b.shape, dtype=dtype or b.dtype, space=space)
        copy_array(out, b)
        return out
    if isinstance(shape, (int, np.integer)):
        shape = (int(shape),)
    shape = tuple(int(s) for s in shape)
    if space == "system":
        return np.zeros(shape, dtype=ringshim.np_dtype(dtype)).view(_HostArray)
    t = rt.zeros(shape, dtype, space)
    if isinstance(dtype, str) and dtype == "ci32":
        t = t.reshape(shape + (2,))
    return _DevArray(t, space, dtype if isinstance(dtype, str) else None)


def _raw(x):
    return x.t if isinstance(x, _DevArray) else x


def copy_array(dst, src):
    """``bifrost.ndarray.copy_array`` between any two spaces (enqueued; callers sync as the reference does)."""
    rt.copy_array(_raw(dst), np.asarray(src) if isinstance(src, _HostArray) else _raw(src))
    return dst
