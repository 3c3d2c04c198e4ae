"""CorrOutputFull: turn CorrAcc's accumulated visibility matrix into per-baseline packets.

Mirror of ``lwa352_pipeline.blocks.corr_output_full_block.CorrOutputFull`` (reference: pipeline/lwa352_pipeline/
blocks/corr_output_full_block.py:364-698): same constructor signature (plus a trailing ``gpu``), the same command
keys (``dest_ip``, ``dest_file``, ``dest_port``, ``max_mbps``), one packet per baseline s0 <= s1 in the same
nested-loop order with the same payload bytes.

The reference runs ``bfXgpuReorder`` on the host into a [stand, stand, pol, pol, chan, 2] matrix
(corr_output_full_block.py:669), then slices and transposes every baseline on the host again
(``send_packets_bf`` :497-548, ``send_packets_py`` :440-477).  Here one device kernel (``xeng_reorder_packets``) goes
from the xGPU-ordered buffer straight to the packet payloads -- [nbl][chan][pol][pol][2] for the "COR" format,
[nbl][pol][pol][chan][2] for the plain format -- and one device->host copy lands them in pinned memory, so the
host side only prepends headers.  The UDP/disk transport itself is bifrost's (``UDPTransmit``/``DiskWriter``,
absent here); the stand-in below writes header + payload to a file object or a UDP socket and is deliberately
simple: transport is outside the X-engine hot path.
"""
import json
import os
import socket
import struct
import time

import numpy as np

from .. import ops
from .. import runtime as rt
from .block_base import Block

COR_SYNC_WORD = 0x5CDEC0DE      # LWA packet sync word as bifrost's packet writers emit it [recollection]


def py_header(sync_time, this_gulp_time, bw_hz, sfreq, acc_len, nchan, chan0, npol, s0, s1):
    """Header of the plain format: ``struct.pack(">QQ2d4I", ...) + struct.pack(">2I", s0, s1)``
    (corr_output_full_block.py:444-459)."""
    return struct.pack(">QQ2d4I", sync_time, this_gulp_time, bw_hz, sfreq, acc_len, nchan, chan0, npol) + \
        struct.pack(">2I", s0, s1)


def cor_header(frame_number, tuning, chan0, gain, time_tag, navg, s0, s1):
    """32-byte header of the "COR" format as laid out in the block's docstring (corr_output_full_block.py:216-228):
    sync_word, id (0x02 = COR) + 24-bit frame number, secs_count (carries the tuning word), freq_count, cor_gain,
    time_tag, cor_navg, stand_i, stand_j (1-based), all big-endian."""
    return struct.pack(">IIIhhqihh", COR_SYNC_WORD, (0x02 << 24) | (frame_number & 0xFFFFFF), tuning & 0xFFFFFFFF,
                       chan0, gain, time_tag, navg, s0 + 1, s1 + 1)


class CorrOutputFull(Block):
    def __init__(self, log, iring, guarantee=True, core=-1, nchan=192, npol=2, nstand=352, etcd_client=None,
                 dest_port=10000, checkfile=None, checkfile_acc_len=1, antpol_to_bl=None, bl_is_conj=None,
                 use_cor_fmt=True, nchan_sum=1, pipeline_idx=1, npipeline=1, gpu=-1):
        super(CorrOutputFull, self).__init__(log, iring, None, guarantee, core, etcd_client=etcd_client)
        # expected-output file in make_golden_inputs.py's format: complex128 [time][nchan][nstand][nstand][npol][npol],
        # each time slice integrated over checkfile_acc_len samples (corr_output_full_block.py:155-166, 395-403)
        self.checkfile_acc_len = checkfile_acc_len
        self.checkfile = open(checkfile, 'rb') if checkfile is not None else None
        self.checkfile_nbytes = os.path.getsize(checkfile) if checkfile is not None else 0
        self.nchan_sum, self.pipeline_idx, self.npipeline = nchan_sum, pipeline_idx, npipeline
        self.nchan, self.npol, self.nstand, self.gpu = nchan, npol, nstand, gpu
        self.words_per_chan = (nstand // 2 + 1) * (nstand // 4) * npol * npol * 4
        self.matlen = nchan * self.words_per_chan
        wrapped_idx = ((pipeline_idx - 1) % npipeline) + 1                 # corr_output_full_block.py:379-381
        self.tuning = ((nchan_sum << 16) | (npipeline << 8) | wrapped_idx) & 0x00FFFFFF
        self.igulp_size = self.matlen * 8
        self.nbl = nstand * (nstand + 1) // 2
        self.payload_words = nchan * npol * npol * 2
        self.dump_size = self.nbl * self.payload_words * 4
        self.use_cor_fmt = use_cor_fmt
        self.layout = ops.PAYLOAD_COR if use_cor_fmt else ops.PAYLOAD_PY
        rt.set_device(self.gpu)
        shape = [nstand, nstand, npol, npol]
        self.antpol_to_bl = np.zeros(shape, dtype=np.int32)
        self.bl_is_conj = np.zeros(shape, dtype=np.int32)
        if antpol_to_bl is not None:
            self.antpol_to_bl[...] = antpol_to_bl
        if bl_is_conj is not None:
            self.bl_is_conj[...] = bl_is_conj
        self._a2bl_gpu = rt.empty(shape, 'i32', 'cuda')
        self._conj_gpu = rt.empty(shape, 'i32', 'cuda')
        self._stage_gpu = None                                             # only needed when the ring is in host memory
        self.payload_gpu = rt.empty([self.nbl, self.payload_words], 'i32', 'cuda')
        self.payloads = rt.zeros([self.nbl, self.payload_words], 'i32', 'cuda_host')
        self.output_file = None
        self.sock = None
        self.sink = None               # optional callable(header_bytes, payload ndarray) used instead of a destination
        self.define_command_key('dest_ip', type=str, initial_val='0.0.0.0')
        self.define_command_key('dest_file', type=str, initial_val='')
        self.define_command_key('dest_port', type=int, initial_val=dest_port)
        self.define_command_key('max_mbps', type=int, initial_val=-1)
        self.update_command_vals()

    # ------------------------------------------------------------------------------------------ device side
    def _upload_maps(self):
        rt.copy_array(self._a2bl_gpu, np.ascontiguousarray(self.antpol_to_bl))
        rt.copy_array(self._conj_gpu, np.ascontiguousarray(self.bl_is_conj))

    def make_payloads(self, span_data):
        """xGPU-ordered int32 [2][nchan][words_per_chan] (device, or host memory staged through the device) ->
        ``self.payloads`` (pinned host), one row per baseline."""
        d_in = span_data
        if not getattr(span_data, "is_cuda", False):
            if self._stage_gpu is None:
                self._stage_gpu = rt.empty(self.igulp_size // 4, 'i32', 'cuda')
            rt.copy_array(self._stage_gpu, span_data)
            d_in = self._stage_gpu
        try:
            ops.reorder_packets(d_in, self.payload_gpu, self._a2bl_gpu, self._conj_gpu, self.nstand, self.npol,
                                self.nchan, self.words_per_chan, self.layout)
        except Exception as e:
            self.log.error("xeng_reorder_packets failed: %s" % e)
            raise RuntimeError(str(e))
        rt.copy_array(self.payloads, self.payload_gpu)
        rt.stream_synchronize()
        return self.payloads

    # ------------------------------------------------------------------------------------------ checkfile
    def get_checkfile_corr(self, t):
        """Integration ``t`` of the check file, wrapping at its end (corr_output_full_block.py:421-437)."""
        dim = (self.nchan, self.nstand, self.nstand, self.npol, self.npol)
        nbytes = int(np.prod(dim)) * 16
        self.checkfile.seek((nbytes * t) % self.checkfile_nbytes)
        raw = self.checkfile.read(nbytes)
        if len(raw) != nbytes:
            self.log.error("CORR OUTPUT >> Failed to get correlation matrix from checkfile")
            return np.zeros(dim, dtype=complex)
        return np.frombuffer(raw, dtype=complex).reshape(dim)

    def check_against_file(self, this_gulp_time, upstream_acc_len):
        """Compare the dump just turned into payloads with the check file, baseline by baseline for s1 >= s0 and
        every pol product, real and imaginary parts separately -- the reference's in-pipeline check of the conjugation
        convention (corr_output_full_block.py:550-603), on the per-baseline payloads instead of the square matrix.
        Returns (good, bad) counts of (baseline, pol product, part) rows; also published as stats."""
        assert upstream_acc_len % self.checkfile_acc_len == 0, \
            "CORR OUTPUT >> Testfile acc len not compatible with pipeline acc len"
        assert this_gulp_time % self.checkfile_acc_len == 0, \
            "CORR OUTPUT >> Testfile acc len not compatible with pipeline start time"
        nblocks = upstream_acc_len // self.checkfile_acc_len
        dtest = np.zeros((self.nchan, self.nstand, self.nstand, self.npol, self.npol), dtype=complex)
        for i in range(nblocks):
            dtest += self.get_checkfile_corr(this_gulp_time // self.checkfile_acc_len + i)
        s0, s1 = np.triu_indices(self.nstand)                       # packet order: s0-major, s1 >= s0
        want = np.moveaxis(dtest[:, s0, s1], 0, 1)                  # [baseline][chan][p0][p1]
        got = (self.payloads.numpy() if hasattr(self.payloads, "numpy") else np.asarray(self.payloads))
        if self.use_cor_fmt:
            got = got.reshape(self.nbl, self.nchan, self.npol, self.npol, 2)
        else:                                                        # [baseline][p0][p1][chan][2]
            got = got.reshape(self.nbl, self.npol, self.npol, self.nchan, 2).transpose(0, 3, 1, 2, 4)
        ok_re = (got[..., 0] == want.real).all(axis=1)              # [baseline][p0][p1]
        ok_im = (got[..., 1] == want.imag).all(axis=1)
        good = int(ok_re.sum() + ok_im.sum())
        bad = int(ok_re.size + ok_im.size - good)
        nonzero = int((got[..., 0] != 0).any(axis=1).sum() + (got[..., 1] != 0).any(axis=1).sum())
        for b in np.argwhere(~(ok_re & ok_im))[:8]:
            self.log.error("CORR OUTPUT >> test vector mismatch! [%d, %d, %d, %d]" % (s0[b[0]], s1[b[0]], b[1], b[2]))
        (self.log.error if bad else self.log.info)(
            "CORR OUTPUT >> test vector check complete. Good: %d, Bad: %d, Non-zero: %d, Zero: %d"
            % (good, bad, nonzero, ok_re.size + ok_im.size - nonzero))
        self.update_stats({'checkfile_good': good, 'checkfile_bad': bad})
        return good, bad

    # ------------------------------------------------------------------------------------------ host side
    def _open_destination(self):
        if self.output_file:
            self.output_file.close()
            self.output_file = None
        if self.command_vals['dest_file'] != "":
            self.log.info("CORR OUTPUT >> Trying to open file %s for output" % self.command_vals['dest_file'])
            try:
                self.output_file = open(self.command_vals['dest_file'], "wb")
            except OSError:
                self.log.error("CORR OUTPUT >> Tried to open file %s for output but failed"
                               % self.command_vals['dest_file'])
        elif self.command_vals['dest_ip'] != "0.0.0.0":
            if self.sock is not None:
                self.sock.close()
            self.sock = socket.socket(socket.AF_INET, socket.SOCK_DGRAM)
            self.sock.connect((self.command_vals['dest_ip'], self.command_vals['dest_port']))

    def _emit(self, header, payload):
        if self.sink is not None:
            self.sink(header, payload)
        elif self.output_file is not None:
            self.output_file.write(header)
            self.output_file.write(payload.tobytes())
        elif self.sock is not None:
            self.sock.send(header + payload.tobytes())

    def send_packets(self, ihdr, this_gulp_time, upstream_acc_len):
        """One packet per baseline, s0-major, as send_packets_bf / send_packets_py do, with the max_mbps throttle
        applied every 50 packets (corr_output_full_block.py:497-548)."""
        start = block_start = time.time()
        payloads = self.payloads.numpy() if hasattr(self.payloads, "numpy") else np.asarray(self.payloads)
        pkt_payload_bits = self.payload_words * 4 * 8
        max_mbps = self.command_vals['max_mbps']
        if self.use_cor_fmt:
            samples_per_spectra = int(ihdr['nchan'] * ihdr['fs_hz'] / ihdr['bw_hz'])
            time_tag = this_gulp_time * samples_per_spectra
            navg = upstream_acc_len * samples_per_spectra
        bl = 0
        block_pkts = 0
        for s0 in range(self.nstand):
            for s1 in range(s0, self.nstand):
                if self.use_cor_fmt:
                    hdr = cor_header(self.nbl, self.tuning, ihdr['chan0'], 0, time_tag, navg, s0, s1)
                else:
                    hdr = py_header(ihdr['sync_time'], this_gulp_time, ihdr['bw_hz'], ihdr['sfreq'], upstream_acc_len,
                                    self.nchan, ihdr['chan0'], self.npol, s0, s1)
                self._emit(hdr, payloads[bl])
                bl += 1
                block_pkts += 1
                if max_mbps > 0 and block_pkts > 50:
                    delay = block_pkts * pkt_payload_bits / (1.e6 * max_mbps) - (time.time() - block_start)
                    if delay > 0:
                        time.sleep(delay)
                    block_start, block_pkts = time.time(), 0
        elapsed = max(time.time() - start, 1e-9)
        self.update_stats({'output_gbps': 8 * self.dump_size / elapsed / 1e9})

    def main(self):
        rt.set_core(self.core)
        rt.set_device(self.gpu)
        self.bind_proclog.update({'ncore': 1, 'core0': rt.get_core()})
        prev = time.time()
        for iseq in self.iring.read(guarantee=self.guarantee):
            self.update_pending = True                       # reprocess commands on each new sequence
            ihdr = json.loads(iseq.header.tostring())
            this_gulp_time = ihdr['seq0']
            upstream_acc_len = ihdr['acc_len']
            if 'ant_to_bl_id' in ihdr:
                self.antpol_to_bl[...] = ihdr['ant_to_bl_id']
            if 'bl_is_conj' in ihdr:
                self.bl_is_conj[...] = ihdr['bl_is_conj']
            self._upload_maps()
            for ispan in iseq.read(self.igulp_size):
                if ispan.size < self.igulp_size:
                    continue
                if self.update_pending:
                    self.update_command_vals()
                    self._open_destination()
                self.update_stats({'curr_sample': this_gulp_time})
                t = time.time()
                acquire_time, prev = t - prev, t
                self.make_payloads(ispan.data)
                if self.checkfile is not None:
                    self.check_against_file(this_gulp_time, upstream_acc_len)
                if self.sink is not None or self.output_file is not None or self.sock is not None:
                    self.send_packets(ihdr, this_gulp_time, upstream_acc_len)
                t = time.time()
                process_time, prev = t - prev, t
                self.perf_proclog.update({'acquire_time': acquire_time, 'reserve_time': 0,
                                          'process_time': process_time})
                self.update_stats()
                this_gulp_time += upstream_acc_len
        if self.output_file:
            self.output_file.close()
        if self.checkfile:
            self.checkfile.close()
