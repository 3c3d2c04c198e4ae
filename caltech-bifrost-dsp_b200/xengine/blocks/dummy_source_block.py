"""DummySource: synthetic 4+4-bit voltages into a ring, with the sequence header the X-engine blocks expect.

Drop-in for ``lwa352_pipeline.blocks.dummy_source_block.DummySource`` (reference: pipeline/lwa352_pipeline/blocks/
dummy_source_block.py:45-330): same constructor signature, same header keys and values (``nchan``, ``system_nchan =
32 nchan``, ``chan0 = 0``, ``bw_hz = 24e3 nchan``, ``fs_hz = 196608000``, ``sfreq = 0``, ``nstand``, ``npol``,
``seq0 = 0``, identity ``input_to_ant`` / ``ant_to_input``, ``sync_time``, ``pipeline_id``; :231-245), same data:
the ramp ``byte = stand % 8`` for every time, channel and polarisation (:194-195), or gulps read round-robin from
``testfile`` (raw ``time x chan x stand x pol`` bytes, :205-219), or nothing at all with ``skip_write``.

One addition at the END of the signature: ``ngulp=None`` stops after that many gulps (the reference runs until
``shutdown()``); with a device ring the gulp is written on the device (an on-device source).
"""
import json
import os
import threading
import time

import numpy as np

from .. import runtime as rt

NTEST_BLOCKS = 4


class DummySource(object):
    def __init__(self, log, oring, ntime_gulp=2500, core=-1, nchan=192, nstand=352, npol=2, skip_write=False,
                 target_throughput=22.0, testfile=None, pipeline_id=0, header={}, ngulp=None):
        self.log, self.oring = log, oring
        self.ntime_gulp, self.pipeline_id, self.core = ntime_gulp, pipeline_id, core
        self.nchan, self.npol, self.nstand = nchan, npol, nstand
        self.ninputs = nstand * npol
        self.skip_write, self.target_throughput = skip_write, target_throughput
        self.header_base = dict(header)
        self.ngulp = ngulp
        cls = type(self).__name__
        for short in ("bind", "in", "out", "size", "sequence0", "perf"):
            setattr(self, ("sequence" if short == "sequence0" else short) + "_proclog", rt.ProcLog("%s/%s" % (cls, short)))
        self.out_proclog.update({'nring': 1, 'ring0': getattr(self.oring, 'name', None)})
        self.size_proclog.update({'nseq_per_gulp': self.ntime_gulp})
        self.gulp_size = ntime_gulp * nchan * nstand * npol            # one byte per 4+4-bit sample
        self.testfile = open(testfile, 'rb') if testfile is not None else None
        self.testfile_nbytes = os.path.getsize(testfile) if testfile is not None else 0
        # input i is (stand, pol) = divmod(i, npol): the identity maps of dummy_source_block.py:176-186
        self.input_to_ant = np.array([[i // npol, i % npol] for i in range(self.ninputs)], dtype=np.int32)
        self.ant_to_input = np.arange(self.ninputs, dtype=np.int32).reshape(nstand, npol)
        self.test_data = np.zeros([ntime_gulp, nchan, nstand, npol], dtype=np.uint8)
        if not skip_write:
            self.test_data[...] = (np.arange(nstand, dtype=np.uint8) % 8)[None, None, :, None]
        self._device_ramp = None
        self.shutdown_event = threading.Event()

    def get_test_data(self):
        """The ramp as complex numbers, as the reference decodes it for its own checks (:200-203)."""
        return (self.test_data >> 4) + 1j * (self.test_data & 0xf)

    def shutdown(self):
        self.shutdown_event.set()

    def get_testfile_gulp(self, t):
        self.testfile.seek((self.gulp_size * t) % self.testfile_nbytes)
        raw = self.testfile.read(self.gulp_size)
        if len(raw) != self.gulp_size:
            self.log.error("Failed to get input test vector gulp")
            return np.zeros_like(self.test_data)
        return np.frombuffer(raw, dtype=np.uint8).reshape(self.test_data.shape)

    def header(self):
        hdr = dict(self.header_base)
        hdr.update({'nchan': self.nchan, 'system_nchan': 32 * self.nchan, 'chan0': 0, 'bw_hz': 24e3 * self.nchan,
                    'fs_hz': 196608000, 'sfreq': 0.0, 'nstand': self.nstand, 'npol': self.npol, 'seq0': 0,
                    'input_to_ant': self.input_to_ant.tolist(), 'ant_to_input': self.ant_to_input.tolist(),
                    'sync_time': int(time.time()), 'pipeline_id': self.pipeline_id})
        return hdr

    def _write(self, ospan, time_tag):
        if self.skip_write:
            return
        if self.testfile:
            rt.copy_array(ospan.data, self.get_testfile_gulp(time_tag))
            return
        if getattr(self.oring, 'space', 'system') == 'cuda':
            if self._device_ramp is None:                  # uploaded once; every gulp is a device-to-device copy
                self._device_ramp = rt.empty(self.gulp_size, 'u8', 'cuda')
                rt.copy_array(self._device_ramp, self.test_data)
            rt.copy_array(ospan.data, self._device_ramp)
        else:
            rt.copy_array(ospan.data, self.test_data)

    def main(self):
        rt.set_core(self.core)
        self.bind_proclog.update({'ncore': 1, 'core0': rt.get_core()})
        self.oring.resize(self.gulp_size, self.gulp_size * 4)
        time_tag, report_bytes, tick = 0, 0, time.time()
        prev = time.time()
        with self.oring.begin_writing() as oring:
            with oring.begin_sequence(time_tag=time_tag, header=json.dumps(self.header())) as oseq:
                while not self.shutdown_event.is_set() and (self.ngulp is None or time_tag < self.ngulp):
                    with oseq.reserve(self.gulp_size) as ospan:
                        t = time.time()
                        reserve_time, prev = t - prev, t
                        self._write(ospan, time_tag)
                        rt.stream_synchronize()            # the span is committed on leaving the with-block
                        time_tag += 1
                    t = time.time()
                    process_time, prev = t - prev, t
                    report_bytes += self.gulp_size
                    gbps = 8 * report_bytes / max(t - tick, 1e-9) / 1e9
                    self.perf_proclog.update({'acquire_time': 0, 'reserve_time': reserve_time,
                                              'process_time': process_time, 'gbps': gbps})
                    if self.target_throughput and gbps > self.target_throughput:     # throttle (:276-283)
                        time.sleep(max(0., 8 * report_bytes / 1e9 / self.target_throughput - (t - tick)))
