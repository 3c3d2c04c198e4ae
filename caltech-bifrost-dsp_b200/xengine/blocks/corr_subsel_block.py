"""CorrSubsel: pick a runtime-selectable set of baselines out of Corr's output and sum adjacent channels.

Drop-in for ``lwa352_pipeline.blocks.corr_subsel_block.CorrSubsel`` (reference: pipeline/lwa352_pipeline/
blocks/corr_subsel_block.py:17-331): same constructor signature, ``baselines`` command key, header keys
(``nchan``, ``nvis``, ``nchan_sum``, ``baselines``, ``sfreq``) and output layout int32 [nchan/nchan_sum][nvis][2]
(re, im).  ``_bf.bfXgpuSubSelect`` (corr_subsel_block.py:298) becomes ``xeng_subsel``.
"""
import json
import time

import numpy as np

from .. import ops
from .. import runtime as rt
from .block_base import Block


class CorrSubsel(Block):
    nvis_out = 48 * 49 * 4 // 2     # 48 stands, dual pol: 4704 (corr_subsel_block.py:185)

    def __init__(self, log, iring, oring, guarantee=True, core=-1, etcd_client=None, nchan=192, npol=2, nstand=352,
                 nchan_sum=4, gpu=-1, antpol_to_bl=None, bl_is_conj=None, payload_layout=None):
        super(CorrSubsel, self).__init__(log, iring, oring, guarantee, core, etcd_client=etcd_client)
        # optional, not in the reference: write the span in CorrOutputPart's packet order ('py' | 'cor') on the
        # device instead of the ring layout, announced by the header key ``payload_layout``
        if payload_layout not in (None, 'py', 'cor'):
            raise ValueError("payload_layout must be None, 'py' or 'cor'")
        self.payload_layout = payload_layout
        self.nchan_in = nchan
        self.nchan_out = nchan // nchan_sum
        self.nchan_sum = nchan_sum
        self.npol, self.nstand, self.gpu = npol, nstand, gpu
        self.words_per_chan = (nstand // 2 + 1) * (nstand // 4) * npol * npol * 4
        self.matlen = self.nchan_in * self.words_per_chan
        rt.set_device(self.gpu)
        self.igulp_size = self.matlen * 8
        self.ogulp_size = self.nchan_out * self.nvis_out * 8

        # selection tables: the host copies are filled by update_subsel(), the device copies are refreshed by
        # main() at sequence boundaries only, so a selection never changes in the middle of a sequence
        self._subsel = rt.empty(self.nvis_out, 'i32', 'cuda')
        self._conj = rt.empty(self.nvis_out, 'i32', 'cuda')
        self._subsel_next = rt.zeros(self.nvis_out, 'i32', 'cuda_host')
        self._conj_next = rt.zeros(self.nvis_out, 'i32', 'cuda_host')
        self.obuf_gpu = rt.empty([self.nchan_out, self.nvis_out], 'ci32', 'cuda')
        self.update_stats()

        shape = [nstand, nstand, npol, npol]
        self._antpol_to_bl = antpol_to_bl if antpol_to_bl is not None else np.zeros(shape, dtype=np.int32)
        self._bl_is_conj = bl_is_conj if bl_is_conj is not None else np.zeros(shape, dtype=np.int32)

        # default selection: pol-0 autocorrelations (corr_subsel_block.py:230)
        subsel = [[[i % nstand, 0], [i % nstand, 0]] for i in range(self.nvis_out)]
        self.define_command_key('baselines', type=list, initial_val=subsel,
                                condition=lambda x: len(x) == self.nvis_out)
        self.update_subsel(subsel)

    def update_subsel(self, baselines):
        """Translate ``baselines`` ([[stand0, pol0], [stand1, pol1]] per output visibility) into buffer word
        indices and conjugation flags.  Takes effect at the next sequence boundary."""
        rt.set_core(self.core)
        bl = np.asarray(baselines, dtype=np.int64).reshape(self.nvis_out, 2, 2)
        s0, p0, s1, p1 = bl[:, 0, 0], bl[:, 0, 1], bl[:, 1, 0], bl[:, 1, 1]
        a2bl, conj = np.asarray(self._antpol_to_bl), np.asarray(self._bl_is_conj)
        rt.copy_array(self._subsel_next, np.ascontiguousarray(a2bl[s0, s1, p0, p1], dtype=np.int32))
        rt.copy_array(self._conj_next, np.ascontiguousarray(conj[s0, s1, p0, p1], dtype=np.int32))

    def _load_selection(self, ohdr):
        self.log.info("Updating baseline subselection indices")
        self.update_command_vals()
        self.update_subsel(self.command_vals['baselines'])
        rt.copy_array(self._subsel, self._subsel_next)
        rt.copy_array(self._conj, self._conj_next)
        ohdr['baselines'] = self.command_vals['baselines']

    def main(self):
        rt.set_core(self.core)
        rt.set_device(self.gpu)
        self.bind_proclog.update({'ncore': 1, 'core0': rt.get_core()})
        self.oring.resize(self.ogulp_size)
        time_tag = 1
        with self.oring.begin_writing() as oring:
            prev = time.time()
            for iseq in self.iring.read(guarantee=self.guarantee):
                ihdr = json.loads(iseq.header.tostring())
                now = ihdr['seq0']
                acc_len = ihdr['acc_len']
                ohdr = dict(ihdr)
                ohdr['nchan'] = ihdr['nchan'] // self.nchan_sum
                ohdr['nvis'] = self.nvis_out
                chan_width = ihdr['bw_hz'] / ihdr['nchan']
                ohdr['sfreq'] = (ihdr['sfreq'] + ((self.nchan_sum - 1) * chan_width)) / self.nchan_sum
                self._load_selection(ohdr)                   # a new sequence always picks up the latest selection
                ohdr['nchan_sum'] = self.nchan_sum
                if self.payload_layout:
                    ohdr['payload_layout'] = self.payload_layout
                oseq = oring.begin_sequence(time_tag=time_tag, header=json.dumps(ohdr), nringlet=iseq.nringlet)
                time_tag += 1
                for ispan in iseq.read(self.igulp_size):
                    if ispan.size < self.igulp_size:
                        continue
                    t = time.time()
                    acquire_time, prev = t - prev, t
                    with oseq.reserve(self.ogulp_size) as ospan:
                        t = time.time()
                        reserve_time, prev = t - prev, t
                        try:
                            if self.payload_layout:
                                ops.subsel_packets(ispan.data, self.obuf_gpu, self._subsel, self._conj, self.nvis_out,
                                                   self.nchan_in, self.nchan_sum, self.words_per_chan,
                                                   ops.PAYLOAD_PY if self.payload_layout == 'py' else ops.PAYLOAD_COR)
                            else:
                                ops.subsel(ispan.data, self.obuf_gpu, self._subsel, self._conj, self.nvis_out,
                                           self.nchan_in, self.nchan_sum, self.words_per_chan)
                        except Exception as e:
                            self.log.error("xeng_subsel failed: %s" % e)
                            raise RuntimeError(str(e))
                        rt.copy_array(ospan.data_view('ci32'), self.obuf_gpu)
                        rt.stream_synchronize()              # the copy must land before the span is committed
                        t = time.time()
                        process_time, prev = t - prev, t
                    self.perf_proclog.update({'acquire_time': acquire_time, 'reserve_time': reserve_time,
                                              'process_time': process_time, 'this_sample': now})
                    now += acc_len
                    if self.update_pending:                  # new selection: new sequence with an updated header
                        oseq.end()
                        self._load_selection(ohdr)
                        ohdr['seq0'] = now
                        oseq = oring.begin_sequence(time_tag=time_tag, header=json.dumps(ohdr),
                                                    nringlet=iseq.nringlet)
                        time_tag += 1
                oseq.end()
