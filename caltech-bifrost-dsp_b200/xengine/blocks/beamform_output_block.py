"""BeamformOutput: packetise the integrated beam powers, one stream per dual-pol beam.

Mirror of ``lwa352_pipeline.blocks.beamform_output_block.BeamformOutput`` (reference: pipeline/lwa352_pipeline/
blocks/beamform_output_block.py:222-386): same constructor signature, list-valued ``dest_ip`` / ``dest_port``
command keys assigned to beams round-robin (:243-245), the same span interpretation -- f32
[nbeam][ntime_gulp][nchan][npol**2] (:353), exactly what ``xeng_bf_integrate`` / ``xeng_bf_run_integrate`` write,
so no device work is needed here -- and one packet per (beam, time step) whose payload is the contiguous
[nchan][4] row ``[XX, YY, re(XY), im(XY)]`` in host byte order (:359).  The header is the one
``send_packets_python`` builds (:311-317): u8 server, beam, tuning, nchan, nbeam, nserver; big-endian u16 navg,
chan0; big-endian u64 seq.  Transport is a stand-in (``sink`` callable or plain UDP sockets): bifrost's
``UDPTransmit('pbeam1_<nchan>')`` is absent and outside the X-engine hot path.
"""
import json
import socket
import time

import numpy as np

from .. import runtime as rt
from .block_base import Block


def pbeam_header(server, beam, tuning, nchan, nbeam, nserver, navg, chan0, seq):
    """beamform_output_block.py:312-314.  The six leading fields are single bytes; values are taken modulo 256, as
    the reference's np.array(..., dtype='>u1') did under the numpy it was written for."""
    return np.array([int(x) & 0xFF for x in (server, beam, tuning, nchan, nbeam, nserver)], dtype='>u1').tobytes() \
        + np.array([int(navg) & 0xFFFF, int(chan0) & 0xFFFF], dtype='>u2').tobytes() + np.array([seq], dtype='>u8').tobytes()


class BeamformOutput(Block):
    def __init__(self, log, iring, guarantee=True, core=-1, etcd_client=None, dest_port=10000, ntime_gulp=480,
                 pipeline_idx=1, nchan=96, nbeam=16):
        super(BeamformOutput, self).__init__(log, iring, None, guarantee, core, etcd_client=etcd_client)
        self.ntime_gulp, self.pipeline_idx, self.nchan, self.nbeam = ntime_gulp, pipeline_idx, nchan, nbeam
        self.define_command_key('dest_ip', type=list, initial_val=['0.0.0.0'])
        self.define_command_key('dest_port', type=list, initial_val=[dest_port])
        self.update_command_vals()
        self.sink = None                     # optional callable(beam, header_bytes, payload_bytes)
        self.socks = [None] * nbeam
        self.beam_ips = [None] * nbeam
        self.beam_ports = [None] * nbeam
        self._assign_destinations()

    def _assign_destinations(self):
        """Round-robin destination lists over the beams; reconnect only the beams that changed (:270-294)."""
        ips, ports = self.command_vals['dest_ip'], self.command_vals['dest_port']
        for beam in range(self.nbeam):
            ip, port = ips[beam % len(ips)], ports[beam % len(ports)]
            if self.beam_ips[beam] == ip and self.beam_ports[beam] == port:
                continue
            self.beam_ips[beam], self.beam_ports[beam] = ip, port
            if self.socks[beam] is not None:
                self.socks[beam].close()
                self.socks[beam] = None
            if ip != '0.0.0.0' and self.sink is None:
                self.log.info("Sending beam %d to %s:%d" % (beam, ip, port))
                self.socks[beam] = socket.socket(socket.AF_INET, socket.SOCK_DGRAM)
                self.socks[beam].connect((ip, port))
        self.update_stats({'dest_ip': self.beam_ips, 'dest_port': self.beam_ports, 'update_pending': False,
                           'last_update_time': time.time()})

    def send_packets(self, server, tuning, nserver, navg, chan0, seq, beam, d):
        """One packet per time step of beam `beam` (:311-317); d is f32 [ntime_gulp][nchan * npol**2]."""
        for t in range(self.ntime_gulp):
            header = pbeam_header(server, beam, tuning, self.nchan, self.nbeam, nserver, navg, chan0, seq)
            if self.sink is not None:
                self.sink(beam, header, d[t].tobytes())
            elif self.socks[beam] is not None:
                self.socks[beam].send(header + d[t].tobytes())
            seq += self.ntime_gulp * navg

    def main(self):
        rt.set_core(self.core)
        self.bind_proclog.update({'ncore': 1, 'core0': rt.get_core()})
        prev = time.time()
        for iseq in self.iring.read(guarantee=self.guarantee):
            ihdr = json.loads(iseq.header.tostring())
            this_gulp_time = ihdr['seq0']
            upstream_acc_len = ihdr['acc_len']
            assert self.nchan == ihdr['nchan']
            assert self.nbeam == ihdr['nbeam']
            nbit, npol, chan0 = ihdr['nbit'], ihdr['npol'], ihdr['chan0']
            npipeline = ihdr['system_nchan'] // self.nchan
            igulp_size = self.ntime_gulp * self.nchan * self.nbeam * npol ** 2 * nbit // 8
            for ispan in iseq.read(igulp_size):
                if ispan.size < igulp_size:
                    continue
                if self.update_pending:
                    self.update_command_vals()
                    self._assign_destinations()
                self.update_stats({'curr_sample': this_gulp_time})
                t = time.time()
                acquire_time, prev = t - prev, t
                data = ispan.data
                if getattr(data, "is_cuda", False):
                    data = data.cpu()
                raw = np.asarray(data.numpy() if hasattr(data, "numpy") else data).reshape(-1).view(np.uint8)
                idata = raw.view(np.float32).reshape(self.nbeam, self.ntime_gulp, self.nchan * npol ** 2)
                for beam in range(self.nbeam):
                    if self.beam_ips[beam] != '0.0.0.0':
                        try:
                            self.send_packets(self.pipeline_idx - 1, 1, npipeline, upstream_acc_len, chan0,
                                              this_gulp_time, beam, idata[beam])
                        except Exception as e:
                            self.log.error("BEAM OUTPUT >> Sending error (beam %d): %s" % (beam, str(e)))
                t = time.time()
                process_time, prev = t - prev, t
                self.perf_proclog.update({'acquire_time': acquire_time, 'reserve_time': 0, 'process_time': process_time})
                self.update_stats({'last_end_sample': this_gulp_time})
                this_gulp_time += upstream_acc_len * self.ntime_gulp
