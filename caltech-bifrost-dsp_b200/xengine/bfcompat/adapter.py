"""XengBF: the object to bind in place of ``bifrost.libbifrost._bf`` for the ten calls the reference's X-engine blocks
make (call sites corr_block.py:253,331,445; corr_subsel_block.py:298; corr_output_full_block.py:669;
beamform_block.py:251,449; beamform_sum_beams_block.py:245; beamform_sum_single_beam_block.py:114).

Same names, same argument order, same return convention (``BF_STATUS_SUCCESS`` or an error code).  Arguments are
whatever ``x.as_BFarray()`` returns: anything with a ``.data`` address (a ctypes BFarray struct from the real
bifrost, or the compat arrays of this package) plus, where a size is needed, ``.shape``.

xGPU is configured at compile time (nstand, nchan, ntime per call are ``#define``s in the reference's build), so
``bfXgpuInitialize(in, out, gpu)`` carries no sizes: give them to the constructor instead.
"""
import ctypes as C

import numpy as np

from .. import _lib, ops

BF_STATUS_SUCCESS = 0
BF_STATUS_INTERNAL_ERROR = 99


def _addr(x):
    d = getattr(x, "data", x)
    if isinstance(d, C.c_void_p):
        return d.value
    if isinstance(d, int):
        return d
    return _lib.ptr(d)


class XengBF(object):
    BF_STATUS_SUCCESS = BF_STATUS_SUCCESS

    def __init__(self, nstand=352, npol=2, nchan=192, ntime_gulp=480, log=None):
        self.cfg = dict(nstand=nstand, npol=npol, nchan=nchan, ntime_gulp=ntime_gulp)
        self.log = log
        self._xc = None
        self._bf = None
        self.last_error = ""

    def _guard(self, fn):
        try:
            fn()
            return BF_STATUS_SUCCESS
        except Exception as e:          # the reference checks the status and raises itself (corr_block.py:254-256)
            self.last_error = str(e)
            if self.log:
                self.log.error("libxeng: %s" % e)
            return BF_STATUS_INTERNAL_ERROR

    # ------------------------------------------------------------------ xGPU wrapper calls
    def bfXgpuInitialize(self, in_arr, out_arr, gpu):
        def go():
            self._xc = ops.XCorr(gpu=gpu, **self.cfg)
        return self._guard(go)

    def bfXgpuKernel(self, in_arr, out_arr, do_dump):
        """Integrate the gulp; on do_dump write the integration to out_arr and clear (xGPU semantics).  The reference
        closes its output span right after a dumping call without a stream sync of its own (corr_block.py:445-450),
        so a dumping call returns only when the device is done."""
        def go():
            self._xc.run(_addr(in_arr), _addr(out_arr), bool(do_dump))
            if do_dump:
                self._xc.sync()
        return self._guard(go)

    def bfXgpuCorrelate(self, in_arr, out_arr, do_dump):
        return self._guard(lambda: self._xc.run_host(_addr(in_arr), _addr(out_arr), bool(do_dump)))

    def bfXgpuGetOrder(self, a2i, a2bl, conj):
        def go():
            lib = _lib.load()
            _lib.check(lib.xeng_get_order(_addr(a2i), _addr(a2bl), _addr(conj), self.cfg["nstand"], self.cfg["npol"]),
                       "xeng_get_order")
        return self._guard(go)

    def bfXgpuSubSelect(self, in_arr, out_arr, vismap, conj, nchan_sum, unused=0):
        def go():
            nvis = int(np.prod(vismap.shape))
            nchan = self.cfg["nchan"]
            ops.subsel(_addr(in_arr), _addr(out_arr), _addr(vismap), _addr(conj), nvis, nchan, int(nchan_sum),
                       ops.words_per_chan(self.cfg["nstand"]))
        return self._guard(go)

    def bfXgpuReorder(self, in_arr, out_arr, a2bl, conj):
        """The reference runs this on HOST arrays (corr_output_full_block.py:669): stage through the device."""
        def go():
            import torch
            ns, npol, nchan = self.cfg["nstand"], self.cfg["npol"], self.cfg["nchan"]
            wpc = ops.words_per_chan(ns)
            nin, nout = 2 * nchan * wpc, ns * ns * npol * npol * nchan * 2

            def host(x, n):
                return torch.from_numpy(np.ctypeslib.as_array((C.c_int32 * n).from_address(_addr(x))))
            d_in = host(in_arr, nin).cuda()
            d_out = torch.empty(nout, dtype=torch.int32, device="cuda")
            ops.reorder(d_in, d_out, host(a2bl, ns * ns * npol * npol).cuda(), host(conj, ns * ns * npol * npol).cuda(),
                        ns, npol, nchan, wpc)
            host(out_arr, nout).copy_(d_out.cpu())
        return self._guard(go)

    # ------------------------------------------------------------------ beamformer library calls
    def bfBeamformInitialize(self, gpu, ninput, nchan, ntime, nbeam, ntime_blocks):
        def go():
            self._bf = ops.Beamformer(gpu=gpu, ninput=ninput, nchan=nchan, ntime=ntime, nbeam=nbeam,
                                      ntime_blocks=ntime_blocks)
            self._ntime_blocks = ntime_blocks
        return self._guard(go)

    def bfBeamformRun(self, in_arr, out_arr, weights):
        def go():
            if getattr(self, "_ntime_blocks", 0):       # Beamform's ntime_sum mode: integrated powers come out
                self._bf.run_integrate(_addr(in_arr), _addr(out_arr), _addr(weights), self._bf.ntime // self._ntime_blocks)
            else:
                self._bf.run(_addr(in_arr), _addr(out_arr), _addr(weights))
        return self._guard(go)

    def bfBeamformIntegrate(self, in_arr, out_arr, ntime_sum):
        return self._guard(lambda: self._bf.integrate(_addr(in_arr), _addr(out_arr), int(ntime_sum)))

    def bfBeamformIntegrateSingleBeam(self, in_arr, out_arr, ntime_sum, beam_id):
        return self._guard(lambda: self._bf.integrate_single(_addr(in_arr), _addr(out_arr), int(ntime_sum), int(beam_id)))

    def sync(self):
        if self._xc is not None:
            self._xc.sync()
        if self._bf is not None:
            self._bf.sync()
