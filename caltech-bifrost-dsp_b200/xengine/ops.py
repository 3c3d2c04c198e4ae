"""Python faces of the libxeng entry points, one per `_bf.*` call the reference blocks make.

Buffers are anything `_lib.ptr` understands (torch CUDA tensors in this repo's tests and bench, bifrost
ndarrays in a bifrost pipeline).  Nothing here computes on the CPU.
"""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import check, ptr


def words_per_chan(nstand, npol=2):
    """corr_block.py:231 with nchan = 1."""
    return (nstand // 2 + 1) * (nstand // 4) * npol * npol * 4


class XCorr(object):
    """xGPU stand-in: bfXgpuInitialize (corr_block.py:253) + bfXgpuKernel (corr_block.py:445)."""

    def __init__(self, nstand=352, npol=2, nchan=192, ntime_gulp=480, gpu=-1, stream=None):
        self._lib = _lib.load()
        self._h = C.c_void_p()
        check(self._lib.xeng_corr_create(C.byref(self._h), nstand, npol, nchan, ntime_gulp, gpu), "xeng_corr_create")
        self.nstand, self.npol, self.nchan, self.ntime_gulp = nstand, npol, nchan, ntime_gulp
        self.matlen = int(self._lib.xeng_corr_matlen(self._h))
        if stream is not None:
            self.set_stream(stream)

    def set_stream(self, stream):
        check(self._lib.xeng_corr_set_stream(self._h, ptr(stream)), "xeng_corr_set_stream")

    def set_impl(self, impl):
        check(self._lib.xeng_corr_set_impl(self._h, int(impl)), "xeng_corr_set_impl")

    def run(self, d_in, d_out, do_dump):
        """bfXgpuKernel(in, out, doDump)."""
        check(self._lib.xeng_corr_run(self._h, ptr(d_in), ptr(d_out), int(bool(do_dump))), "xeng_corr_run")

    def run_host(self, h_in, h_out, do_dump):
        """bfXgpuCorrelate(in, out, doDump): host buffers (xgpu_test.py:86-89)."""
        check(self._lib.xeng_corr_run_host(self._h, ptr(h_in), ptr(h_out), int(bool(do_dump))), "xeng_corr_run_host")

    def accumulate(self, d_in, d_acc, first):
        """Fused Corr + CorrAcc step: d_acc (=|+=) X(d_in)."""
        check(self._lib.xeng_corr_accumulate(self._h, ptr(d_in), ptr(d_acc), int(bool(first))), "xeng_corr_accumulate")

    def set_subsel(self, vismap, conj, nchan_sum=4):
        """Selection for the fused fast-dump path: the host arrays CorrSubsel.update_subsel builds
        (corr_subsel_block.py:237-250).  An empty selection clears it."""
        vm = np.ascontiguousarray(vismap, dtype=np.int32)
        cj = np.ascontiguousarray(conj, dtype=np.int32)
        if vm.shape != cj.shape or vm.ndim != 1:
            raise ValueError("vismap and conj must be 1-D arrays of the same length")
        check(self._lib.xeng_corr_set_subsel(self._h, ptr(vm), ptr(cj), int(vm.size), int(nchan_sum)),
              "xeng_corr_set_subsel")
        self.nvis, self.nchan_sum = int(vm.size), int(nchan_sum)

    def accumulate_subsel(self, d_in, d_acc, first, d_sub_out, sub_first=True):
        """Fused Corr + CorrAcc + CorrSubsel step: d_acc (=|+=) X(d_in); d_sub_out[nchan/nchan_sum][nvis][2] (=|+=) this
        gulp's selected visibilities summed over nchan_sum channels."""
        check(self._lib.xeng_corr_accumulate_subsel(self._h, ptr(d_in), ptr(d_acc), int(bool(first)), ptr(d_sub_out),
                                                    int(bool(sub_first))), "xeng_corr_accumulate_subsel")

    def sync(self):
        check(self._lib.xeng_corr_sync(self._h), "xeng_corr_sync")

    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            self._lib.xeng_corr_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def get_order(ant_to_input):
    """bfXgpuGetOrder (corr_block.py:331-333): returns (antpol_to_bl, bl_is_conj) int32 [ns,ns,np,np]."""
    a2i = np.ascontiguousarray(ant_to_input, dtype=np.int32)
    nstand, npol = a2i.shape
    a2bl = np.zeros([nstand, nstand, npol, npol], dtype=np.int32)
    conj = np.zeros_like(a2bl)
    check(_lib.load().xeng_get_order(ptr(a2i), ptr(a2bl), ptr(conj), nstand, npol), "xeng_get_order")
    return a2bl, conj


def acc(d_acc, d_in, n, first, stream=None):
    """BFMap 'a = b' / 'a += b' (corr_acc_block.py:299-306) over n int32 words."""
    check(_lib.load().xeng_acc(ptr(d_acc), ptr(d_in), int(n), int(bool(first)), ptr(stream)), "xeng_acc")


def subsel(d_in, d_out, d_vismap, d_conj, nvis, nchan, nchan_sum, wpc, stream=None):
    """bfXgpuSubSelect (corr_subsel_block.py:298)."""
    check(_lib.load().xeng_subsel(ptr(d_in), ptr(d_out), ptr(d_vismap), ptr(d_conj), nvis, nchan, nchan_sum, wpc,
                                  ptr(stream)), "xeng_subsel")


def reorder(d_in, d_out, d_a2bl, d_conj, nstand, npol, nchan, wpc, stream=None):
    """bfXgpuReorder (corr_output_full_block.py:669), on the device."""
    check(_lib.load().xeng_reorder(ptr(d_in), ptr(d_out), ptr(d_a2bl), ptr(d_conj), nstand, npol, nchan, wpc,
                                   ptr(stream)), "xeng_reorder")


PAYLOAD_COR, PAYLOAD_PY = 1, 2
SNAP2_HEADER_BYTES = 32


def snap2_assemble(d_pkts, npkt, pkt_stride, d_out, seq0, ntime, nchan, ninput, first_chan=0, zero_first=True,
                   d_nvalid=None, stream=None):
    """SNAP2 packets (device copy of the raw receive buffer) -> u8 [ntime][nchan][ninput] voltages; the host-side
    counterpart in the reference is bifrost's UDP capture behind Capture (capture_block.py:219-305)."""
    check(_lib.load().xeng_snap2_assemble(ptr(d_pkts), int(npkt), int(pkt_stride), ptr(d_out), int(seq0), int(ntime),
                                          int(nchan), int(ninput), int(first_chan), int(bool(zero_first)),
                                          ptr(d_nvalid), ptr(stream)), "xeng_snap2_assemble")


UPCHAN_FINE_LAST, UPCHAN_FREQ_MERGED = 0, 1


def upchan_fft(d_in, d_out, nframe, nfine, nchan, ninput, layout=UPCHAN_FINE_LAST, stream=None):
    """transpose -> fft(axes='fine_time') [-> transpose] of the offline scripts as one kernel
    (lwa352-upchan-bf.py:94-96, lwa352-upchan-imag.py:93-98).  d_in u8 [nframe][nfine][nchan][ninput];
    d_out cf32 [nframe][nchan][ninput][nfine] (FINE_LAST) or [nframe][nchan][nfine][ninput] (FREQ_MERGED)."""
    check(_lib.load().xeng_upchan_fft(ptr(d_in), ptr(d_out), int(nframe), int(nfine), int(nchan), int(ninput),
                                      int(layout), ptr(stream)), "xeng_upchan_fft")


def upchan_beamform(d_in, d_weights, d_out, nframe, nfine, nchan, ninput, nbeam, stream=None):
    """lwa352-upchan-bf.py:94-113 (upchannelise, BfOfflineWeightsBlock's multiply, reduce over stand and pol) as
    one kernel.  d_weights cf32 [nbeam][nchan][ninput][nfine]; d_out cf32 [nframe][nbeam][nchan][nfine]."""
    check(_lib.load().xeng_upchan_beamform(ptr(d_in), ptr(d_weights), ptr(d_out), int(nframe), int(nfine), int(nchan),
                                           int(ninput), int(nbeam), ptr(stream)), "xeng_upchan_beamform")


def upchan_correlate(d_in, d_out, nframe, nfreq, ninput, first=True, stream=None):
    """lwa352-upchan-imag.py:93-106's blocks.correlate on the upchannelised voltages: d_out[f][i][j] (=|+=)
    sum_t d_in[t][f][i] * conj(d_in[t][f][j]) for the 64 x 64 tiles on or below the diagonal.  cf32 device buffers:
    d_in [nframe][nfreq][ninput], d_out [nfreq][ninput][ninput]."""
    check(_lib.load().xeng_upchan_correlate(ptr(d_in), ptr(d_out), int(nframe), int(nfreq), int(ninput),
                                            int(bool(first)), ptr(stream)), "xeng_upchan_correlate")


def subsel_packets(d_in, d_out, d_vismap, d_conj, nvis, nchan, nchan_sum, wpc, layout=PAYLOAD_PY, stream=None):
    """bfXgpuSubSelect written in CorrOutputPart's packet order (corr_output_part_block.py:378-391, 429-463):
    COR = int32 [nvis/4][nchan/nchan_sum][4][2]; PY = big-endian int32 [nvis][nchan/nchan_sum][2]."""
    check(_lib.load().xeng_subsel_packets(ptr(d_in), ptr(d_out), ptr(d_vismap), ptr(d_conj), nvis, nchan, nchan_sum,
                                          wpc, layout, ptr(stream)), "xeng_subsel_packets")


def reorder_packets(d_in, d_out, d_a2bl, d_conj, nstand, npol, nchan, wpc, layout=PAYLOAD_COR, stream=None):
    """bfXgpuReorder fused with the per-baseline payload transpose of CorrOutputFull.send_packets_bf / _py
    (corr_output_full_block.py:505-511, 457-460): int32 [nbl][nchan][npol][npol][2] (COR) or
    [nbl][npol][npol][nchan][2] (PY), baselines s0 <= s1, s0-major."""
    check(_lib.load().xeng_reorder_packets(ptr(d_in), ptr(d_out), ptr(d_a2bl), ptr(d_conj), nstand, npol, nchan, wpc,
                                           layout, ptr(stream)), "xeng_reorder_packets")


def memcpy_async(dst, src, nbytes, stream=None):
    """cudaMemcpyAsync(cudaMemcpyDefault) on `stream`: a copy-engine copy between any two unified addresses (device,
    pinned host, peer memory mapped into this process)."""
    check(_lib.load().xeng_memcpy_async(ptr(dst), ptr(src), int(nbytes), ptr(stream)), "xeng_memcpy_async")


def peer_move(dst, src, nbytes, dst_seq=None, seq=0, stream=None):
    """peer_push that leaves `src` zeroed: the product is moved to its receiver (include/xeng.h, xeng_peer_move)."""
    check(_lib.load().xeng_peer_move(ptr(dst), ptr(src), int(nbytes), ptr(dst_seq), int(seq), ptr(stream)),
          "xeng_peer_move")


def peer_push(dst, src, nbytes, dst_seq=None, seq=0, stream=None):
    """`nbytes` from src to dst by SM stores, then (dst_seq given) the 64-bit word `seq` to dst_seq, released after the
    payload.  dst / dst_seq may be peer memory mapped into this process.  See include/xeng.h (xeng_peer_push)."""
    check(_lib.load().xeng_peer_push(ptr(dst), ptr(src), int(nbytes), ptr(dst_seq), int(seq), ptr(stream)),
          "xeng_peer_push")


class Beamformer(object):
    """bfBeamformInitialize / Run / Integrate / IntegrateSingleBeam
    (beamform_block.py:251,449; beamform_sum_beams_block.py:245; beamform_sum_single_beam_block.py:114)."""

    def __init__(self, gpu=-1, ninput=704, nchan=192, ntime=480, nbeam=16, ntime_blocks=0, stream=None):
        self._lib = _lib.load()
        self._h = C.c_void_p()
        check(self._lib.xeng_bf_create(C.byref(self._h), gpu, ninput, nchan, ntime, nbeam, ntime_blocks), "xeng_bf_create")
        self.ninput, self.nchan, self.ntime, self.nbeam = ninput, nchan, ntime, nbeam
        if stream is not None:
            self.set_stream(stream)

    def set_stream(self, stream):
        check(self._lib.xeng_bf_set_stream(self._h, ptr(stream)), "xeng_bf_set_stream")

    def run(self, d_in, d_out, d_weights=None):
        """bfBeamformRun(in, out, weights).  d_weights=None: the weights last given to set_weights."""
        check(self._lib.xeng_bf_run(self._h, ptr(d_in), ptr(d_out), ptr(d_weights)), "xeng_bf_run")

    def set_weights(self, d_weights):
        """Convert cf32 [nchan][nbeam][ninput] weights to the tensor-core kernel's int8 digit planes once; later
        run / run_integrate calls with d_weights=None use them.  The buffer must outlive its use."""
        check(self._lib.xeng_bf_set_weights(self._h, ptr(d_weights)), "xeng_bf_set_weights")

    def set_impl(self, impl):
        """0: tcgen05 int8 kernel when the shape allows (default); 1: fp32 FMA kernel (cross-check)."""
        check(self._lib.xeng_bf_set_impl(self._h, int(impl)), "xeng_bf_set_impl")

    def set_tap(self, on):
        """Keep the voltage beams of run_integrate in the handle (two kernels) instead of the fused power epilogue."""
        check(self._lib.xeng_bf_set_tap(self._h, int(bool(on))), "xeng_bf_set_tap")

    def run_integrate(self, d_in, d_out, d_weights, ntime_sum):    # d_weights may be None (set_weights)
        """Beamform's ntime_sum mode: integrated powers go to d_out; one fused kernel for ntime_sum <= 128."""
        check(self._lib.xeng_bf_run_integrate(self._h, ptr(d_in), ptr(d_out), ptr(d_weights), int(ntime_sum)),
              "xeng_bf_run_integrate")

    def integrate(self, d_in, d_out, ntime_sum):
        check(self._lib.xeng_bf_integrate(self._h, ptr(d_in), ptr(d_out), int(ntime_sum)), "xeng_bf_integrate")

    def integrate_single(self, d_in, d_out, ntime_sum, beam_id):
        check(self._lib.xeng_bf_integrate_single(self._h, ptr(d_in), ptr(d_out), int(ntime_sum), int(beam_id)),
              "xeng_bf_integrate_single")

    def vlbi_tap(self, d_in, d_out, nbeam_send):
        """First nbeam_send voltage beams as [ntime][nchan][nbeam_send] (beamform_vlbi_output_block.py:260-264)."""
        check(self._lib.xeng_bf_vlbi_tap(self._h, ptr(d_in), ptr(d_out), int(nbeam_send)), "xeng_bf_vlbi_tap")

    def sync(self):
        check(self._lib.xeng_bf_sync(self._h), "xeng_bf_sync")

    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            self._lib.xeng_bf_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
