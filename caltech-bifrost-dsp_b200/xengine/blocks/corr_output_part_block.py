"""CorrOutputPart: packetise CorrSubsel's fast visibilities.

Mirror of ``lwa352_pipeline.blocks.corr_output_part_block.CorrOutputPart`` (reference: pipeline/lwa352_pipeline/
blocks/corr_output_part_block.py:318-475): same constructor signature and command keys (``dest_ip``,
``dest_port``), the same two packet formats with the same payload bytes:

* plain (``use_cor_fmt=False``): ``nvis_per_packet`` visibilities per packet, payload big-endian int32
  [nvis_per_packet][nchan][2], header ``struct.pack(">QQ2d4I", ...)`` + the packet's baselines as big-endian int32
  (:346-364);
* "COR" (``use_cor_fmt=True``): one dual-pol baseline (4 visibilities) per packet, payload int32 [nchan][2][2][2],
  baselines enumerated stand-by-stand over ``nstand_virt`` virtual stands (:366-401).

The reference transposes every span on the host ([chan, vis, 2] -> [vis, chan, 2]).  When the upstream CorrSubsel
was built with ``payload_layout='py'`` / ``'cor'`` the span already IS the payload array (written by
``xeng_subsel_packets`` on the device; header key ``payload_layout``) and this block only slices it; a span in the
reference's ring layout is still accepted and transposed here, as the reference does.  Transport is a stand-in
(``sink`` callable or a UDP socket): bifrost's ``UDPTransmit`` is absent and outside the X-engine hot path.
"""
import json
import socket
import struct
import time

import numpy as np

from .. import runtime as rt
from .block_base import Block
from .corr_output_full_block import cor_header

COR_NPOL = 2


class CorrOutputPart(Block):
    def __init__(self, log, iring, use_cor_fmt=False, guarantee=True, core=-1, etcd_client=None, dest_port=10001,
                 nvis_per_packet=16, nchan_sum=1, pipeline_idx=1, npipeline=1):
        super(CorrOutputPart, self).__init__(log, iring, None, guarantee, core, etcd_client=etcd_client)
        self.nvis_per_packet = nvis_per_packet
        self.nchan_sum, self.pipeline_idx, self.npipeline = nchan_sum, pipeline_idx, npipeline
        wrapped_idx = ((pipeline_idx - 1) % npipeline) + 1                 # corr_output_part_block.py:330-333
        self.tuning = ((nchan_sum << 16) | (npipeline << 8) | wrapped_idx) & 0x00FFFFFF
        self.use_cor_fmt = use_cor_fmt
        self.sock = None
        self.sink = None               # optional callable(header_bytes, payload_bytes) used instead of the socket
        self.define_command_key('dest_ip', type=str, initial_val='0.0.0.0')
        self.define_command_key('dest_port', type=int, initial_val=dest_port)
        self.update_command_vals()

    def _emit(self, header, payload):
        if self.sink is not None:
            self.sink(header, payload)
        elif self.sock is not None:
            self.sock.send(header + payload)

    def _payload_array(self, span, ihdr, nchan, nvis):
        """The span as this block's payload array: a view when CorrSubsel already wrote packet order, else the
        reference's host transpose."""
        raw = np.asarray(span.numpy() if hasattr(span, "numpy") else span).reshape(-1).view(np.uint8)
        want = 'cor' if self.use_cor_fmt else 'py'
        have = ihdr.get('payload_layout')
        if have == want:
            if want == 'py':
                return raw.view('>i4').reshape(nvis, nchan, 2)
            return raw.view(np.int32).reshape(nvis // 4, nchan, 4, 2)
        if have is not None:
            raise RuntimeError("CORR PART OUTPUT >> upstream payload_layout %r does not match this block's format" % have)
        idata = raw.view(np.int32).reshape(nchan, nvis, 2)
        if want == 'py':
            dout = np.zeros([nvis, nchan, 2], dtype='>i4')                # corr_output_part_block.py:429,460-463
            dout[...] = idata.transpose([1, 0, 2])
            return dout
        dview = idata.reshape(nchan, nvis // 4, 4, 2)                      # corr_output_part_block.py:378-391
        return np.ascontiguousarray(dview.transpose([1, 0, 2, 3]))

    def send_packets(self, payloads, baselines, ihdr, this_gulp_time, upstream_acc_len, nchan, nvis):
        start = time.time()
        if not self.use_cor_fmt:
            flat = baselines.flatten()
            n = self.nvis_per_packet
            for vn in range(len(baselines) // n):
                header = struct.pack(">QQ2d4I", ihdr['sync_time'], this_gulp_time, ihdr['bw_hz'], ihdr['sfreq'],
                                     upstream_acc_len, n, nchan, ihdr['chan0']) + flat[vn * 4 * n:(vn + 1) * 4 * n].tobytes()
                self._emit(header, payloads[vn * n:(vn + 1) * n].tobytes())
        else:
            spectra = int(ihdr['nchan_sum'] * nchan * ihdr['fs_hz'] / ihdr['bw_hz'])
            time_tag, navg = this_gulp_time * spectra, upstream_acc_len * spectra
            nbl = nvis // (COR_NPOL ** 2)
            nstand_virt = int((-1 + np.sqrt(1 + 2 * nvis)) / 2)          # effective number of stands (:376)
            bl = 0
            for i in range(nstand_virt):
                for j in range(i, nstand_virt):
                    self._emit(cor_header(nbl, self.tuning, ihdr['chan0'], 0, time_tag, navg, i, j), payloads[bl].tobytes())
                    bl += 1
            assert bl == nbl
        elapsed = max(time.time() - start, 1e-9)
        self.update_stats({'output_gbps': 8 * 8 * nvis * nchan / elapsed / 1e9})

    def main(self):
        rt.set_core(self.core)
        self.bind_proclog.update({'ncore': 1, 'core0': rt.get_core()})
        prev = time.time()
        for iseq in self.iring.read(guarantee=self.guarantee):
            self.update_pending = True
            ihdr = json.loads(iseq.header.tostring())
            this_gulp_time = ihdr['seq0']
            upstream_acc_len = ihdr['acc_len']
            baselines = np.array(ihdr['baselines'], dtype='>i4')
            nchan, nvis = ihdr['nchan'], ihdr['nvis']
            igulp_size = nvis * nchan * 8
            for ispan in iseq.read(igulp_size):
                if ispan.size < igulp_size:
                    continue
                if self.update_pending:
                    self.update_command_vals()
                    self.log.info("CORR PART OUTPUT >> Updating destination to %s:%s"
                                  % (self.command_vals['dest_ip'], self.command_vals['dest_port']))
                    if self.sock is not None:
                        self.sock.close()
                        self.sock = None
                    if self.command_vals['dest_ip'] != "0.0.0.0":
                        self.sock = socket.socket(socket.AF_INET, socket.SOCK_DGRAM)
                        self.sock.connect((self.command_vals['dest_ip'], self.command_vals['dest_port']))
                self.update_stats({'curr_sample': this_gulp_time})
                t = time.time()
                acquire_time, prev = t - prev, t
                if self.sink is not None or self.sock is not None:
                    data = ispan.data
                    if getattr(data, "is_cuda", False):
                        data = data.cpu()
                    payloads = self._payload_array(data, ihdr, nchan, nvis)
                    self.send_packets(payloads, baselines, ihdr, this_gulp_time, upstream_acc_len, nchan, nvis)
                t = time.time()
                process_time, prev = t - prev, t
                self.perf_proclog.update({'acquire_time': acquire_time, 'reserve_time': 0, 'process_time': process_time})
                this_gulp_time += upstream_acc_len
