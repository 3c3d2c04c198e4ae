"""Test doubles that let the blocks' HOST logic (state machines, headers, command handling) run without a GPU.

They are installed by the ``host_only`` fixture (conftest.py) in place of ``xengine.ops`` entry points and of
device allocation.  They do no X-engine arithmetic: each "kernel" just moves a tag (the first int32 / first byte of
its input) so that a test can tell WHICH inputs were combined into WHICH output span.  The real kernels are
covered by the ``-m gpu`` tests.
"""
import numpy as np

CALLS = []


def _np(buf):
    """uint8 numpy view of a span buffer (torch CPU tensor or numpy array)."""
    buf = getattr(buf, "_xeng_raw", buf)       # array of a foreign ring type wrapping a tensor / ndarray
    if hasattr(buf, "numpy"):
        buf = buf.numpy()
    return buf.reshape(-1).view(np.uint8)


class FakeXCorr(object):
    """accumulate(): out.int32[0] (=|+=) in.uint8[0]; out.int32[1] counts the calls of the integration."""

    def __init__(self, nstand=352, npol=2, nchan=192, ntime_gulp=480, gpu=-1, stream=None):
        self.matlen = nchan * (nstand // 2 + 1) * (nstand // 4) * npol * npol * 4
        self.impl = 0

    def set_impl(self, impl):
        self.impl = impl

    def accumulate(self, d_in, d_acc, first):
        tag = int(_np(d_in)[0])
        out = _np(d_acc).view(np.int32)
        if first:
            out[0], out[1] = tag, 1
        else:
            out[0] += tag
            out[1] += 1
        CALLS.append(("corr", tag, bool(first)))

    def set_subsel(self, vismap, conj, nchan_sum=4):
        self.sel0 = int(np.asarray(vismap).reshape(-1)[0]) if np.asarray(vismap).size else -1
        CALLS.append(("set_subsel", self.sel0, int(nchan_sum)))

    def accumulate_subsel(self, d_in, d_acc, first, d_sub_out, sub_first=True):
        """accumulate() + the fast dump: sub.int32[0] (=|+=) tag, sub.int32[1] = first entry of the selection."""
        self.accumulate(d_in, d_acc, first)
        tag = int(_np(d_in)[0])
        sub = _np(d_sub_out).view(np.int32)
        if sub_first:
            sub[0] = tag
        else:
            sub[0] += tag
        sub[1] = getattr(self, "sel0", -1)
        CALLS.append(("corr_sub", tag, bool(first), bool(sub_first)))

    def sync(self):
        CALLS.append(("corr_sync",))

    def close(self):
        pass


def fake_acc(d_acc, d_in, n, first, stream=None):
    a, b = _np(d_acc).view(np.int32), _np(d_in).view(np.int32)
    if first:
        a[:2] = b[:2]
    else:
        a[:2] += b[:2]
    CALLS.append(("acc", int(b[0]), bool(first)))


def fake_subsel(d_in, d_out, d_vismap, d_conj, nvis, nchan, nchan_sum, wpc, stream=None):
    i, o = _np(d_in).view(np.int32), _np(d_out).view(np.int32)
    vm = _np(d_vismap).view(np.int32)
    o[0] = i[0]                 # tag of the integration
    o[1] = vm[0]                # first entry of the selection table in force
    CALLS.append(("subsel", int(i[0]), int(vm[0])))


def fake_reorder_packets(d_in, d_out, d_a2bl, d_conj, nstand, npol, nchan, wpc, layout=1, stream=None):
    i, o = _np(d_in).view(np.int32), _np(d_out).view(np.int32)
    o[0] = i[0]                 # tag of the dump
    o[1] = layout
    CALLS.append(("reorder_packets", int(i[0]), int(layout), int(_np(d_a2bl).view(np.int32)[1])))


class FakeBeamformer(object):
    def __init__(self, gpu=-1, ninput=704, nchan=192, ntime=480, nbeam=16, ntime_blocks=0, stream=None):
        self.dims = (ninput, nchan, ntime, nbeam, ntime_blocks)

    def set_weights(self, d_w):
        self.w = _np(d_w).view(np.complex64).copy()      # a snapshot, as the digit-plane conversion is
        CALLS.append(("bf_set_weights",))

    def run(self, d_in, d_out, d_w=None):
        _np(d_out)[0] = _np(d_in)[0]
        w = _np(d_w).view(np.complex64) if d_w is not None else self.w
        CALLS.append(("bf_run", int(_np(d_in)[0]), w.copy()))

    def run_integrate(self, d_in, d_out, d_w, ntime_sum):
        _np(d_out)[0] = _np(d_in)[0]
        CALLS.append(("bf_run_integrate", int(_np(d_in)[0]), ntime_sum))

    def integrate(self, d_in, d_out, ntime_sum):
        _np(d_out)[0] = _np(d_in)[0]
        CALLS.append(("bf_integrate", int(_np(d_in)[0]), ntime_sum))

    def integrate_single(self, d_in, d_out, ntime_sum, beam_id):
        _np(d_out)[0] = _np(d_in)[0]
        CALLS.append(("bf_integrate_single", int(_np(d_in)[0]), ntime_sum, beam_id))

    def sync(self):
        pass

    def close(self):
        pass
