"""BeamformVlbiOutput: stream the first dual-pol voltage beam, one packet per time sample.

Mirror of ``lwa352_pipeline.blocks.beamform_vlbi_output_block.BeamformVlbiOutput`` (reference: pipeline/
lwa352_pipeline/blocks/beamform_vlbi_output_block.py:188-300): same constructor signature and command keys, span =
cf32 [nchan][nbeam][ntime_gulp] (Beamform's output), of which the first ``(2 // npol) * nbeam_send`` single-pol beams
are sent as one cf32 [nchan][beam] row per time sample (:260-262).  The reference copies the sub-array to the host
and transposes it there twice; here a device span goes through ``xeng_bf_vlbi_tap`` (one kernel writing
[ntime][nchan][nsend]) and only the tap is copied.  Packets are bursts of ``_npacket_burst`` rows throttled to
``_max_bps`` (:263-273).  Transport is a stand-in (``sink`` callable or a UDP socket) for bifrost's
``UDPTransmit('ibeam<nbeam_send>_<nchan>')``; the ibeam header is bifrost's and is not restated: ``sink`` receives
(seq, payload_bytes).
"""
import json
import socket
import time

import numpy as np

from .. import ops
from .. import runtime as rt
from .block_base import Block


class BeamformVlbiOutput(Block):
    def __init__(self, log, iring, guarantee=True, core=-1, etcd_client=None, dest_port=10000, ntime_gulp=480,
                 pipeline_idx=1):
        super(BeamformVlbiOutput, self).__init__(log, iring, None, guarantee, core, etcd_client=etcd_client)
        rt.set_core(self.core)
        self.sock = None
        self.sink = None
        self.define_command_key('dest_ip', type=str, initial_val='0.0.0.0')
        self.define_command_key('dest_port', type=int, initial_val=dest_port)
        self.update_command_vals()
        self.dest_ip, self.dest_port = self.command_vals['dest_ip'], self.command_vals['dest_port']
        self.ntime_gulp = ntime_gulp
        self._npacket_burst = 32
        self._max_bps = 0.6 * 1e9
        self.pipeline_idx = pipeline_idx
        self.nbeam_send = 1
        self.npol = 2     # single-pol upstream beams are read as dual-pol pairs (:206)
        self._bf = None
        self._tap = None

    def _tapped(self, data, nchan, nbeam, nsend):
        """cf32 [ntime][nchan][nsend] on the host: the device kernel for a device span, numpy for a host span."""
        if getattr(data, "is_cuda", False):
            import torch
            if self._bf is None or self._tap is None or self._tap.shape != (self.ntime_gulp, nchan, nsend, 2):
                self._bf = ops.Beamformer(gpu=-1, ninput=2, nchan=nchan, ntime=self.ntime_gulp, nbeam=nbeam)
                self._tap = torch.empty((self.ntime_gulp, nchan, nsend, 2), dtype=torch.float32, device=data.device)
            try:
                self._bf.vlbi_tap(data, self._tap, nsend)
                self._bf.sync()
            except Exception as e:
                self.log.error("xeng_bf_vlbi_tap failed: %s" % e)
                raise RuntimeError(str(e))
            return self._tap.cpu().numpy().view(np.complex64).reshape(self.ntime_gulp, nchan, nsend)
        raw = np.asarray(data.numpy() if hasattr(data, "numpy") else data).reshape(-1).view(np.uint8)
        idata = raw.view(np.complex64).reshape(nchan, nbeam, self.ntime_gulp)
        return np.ascontiguousarray(idata[:, 0:nsend, :].transpose(2, 0, 1))      # :261

    def main(self):
        rt.set_core(self.core)
        self.bind_proclog.update({'ncore': 1, 'core0': rt.get_core()})
        prev = time.time()
        for iseq in self.iring.read(guarantee=self.guarantee):
            self.update_pending = True
            ihdr = json.loads(iseq.header.tostring())
            this_gulp_time = ihdr['seq0']
            nchan, nbeam, nbit, npol = ihdr['nchan'], ihdr['nbeam'], ihdr['nbit'], ihdr['npol']
            igulp_size = self.ntime_gulp * nbeam * nchan * npol * 2 * nbit // 8
            nsend = (self.npol // npol) * self.nbeam_send
            for ispan in iseq.read(igulp_size):
                if ispan.size < igulp_size:
                    continue
                if self.update_pending:
                    self.update_command_vals()
                    if not (self.dest_ip == self.command_vals['dest_ip'] and self.dest_port == self.command_vals['dest_port']):
                        self.dest_ip, self.dest_port = self.command_vals['dest_ip'], self.command_vals['dest_port']
                        self.log.info("VLBI OUTPUT >> Updating destination to %s:%s" % (self.dest_ip, self.dest_port))
                        if self.sock is not None:
                            self.sock.close()
                            self.sock = None
                        if self.sink is None and self.dest_ip != '0.0.0.0':
                            self.sock = socket.socket(socket.AF_INET, socket.SOCK_DGRAM)
                            self.sock.connect((self.dest_ip, self.dest_port))
                    self.update_stats({'dest_ip': self.dest_ip, 'dest_port': self.dest_port,
                                       'update_pending': self.update_pending, 'last_update_time': time.time()})
                self.update_stats({'curr_sample': this_gulp_time})
                t = time.time()
                acquire_time, prev = t - prev, t
                if self.command_vals['dest_ip'] != '0.0.0.0':
                    rows = self._tapped(ispan.data, nchan, nbeam, nsend).reshape(self.ntime_gulp, nchan * nsend)
                    burst_bits = self._npacket_burst * nchan * nsend * 2 * 32
                    toff = 0
                    while toff < self.ntime_gulp:
                        t0 = time.time()
                        for k in range(toff, min(toff + self._npacket_burst, self.ntime_gulp)):
                            if self.sink is not None:
                                self.sink(this_gulp_time + k, rows[k].tobytes())
                            elif self.sock is not None:
                                self.sock.send(rows[k].tobytes())
                        toff += self._npacket_burst
                        delay = burst_bits / self._max_bps - (time.time() - t0)
                        if delay > 0 and self.sink is None:
                            time.sleep(delay)
                t = time.time()
                process_time, prev = t - prev, t
                self.perf_proclog.update({'acquire_time': acquire_time, 'reserve_time': 0, 'process_time': process_time})
                self.update_stats({'last_end_sample': this_gulp_time})
                this_gulp_time += self.ntime_gulp
