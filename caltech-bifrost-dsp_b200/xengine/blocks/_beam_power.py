"""Shared loop of the three beam-power blocks (BeamformSumBeams, BeamformSum, BeamformSumSingleBeam).

All three read cf32 voltage beams [nchan][nbeam][ntime_gulp] and emit, for dual-pol beam p (X = beam 2p,
Y = beam 2p+1) and time block T of ``ntime_sum`` samples, [sum|X|^2, sum|Y|^2, Re sum X Y*, Im sum X Y*] as
f32 [...][T][nchan][4] (reference: beamform_sum_beams_block.py:189-258, beamform_sum_block.py:59-128,
beamform_sum_single_beam_block.py:56-127; arithmetic pinned by verification/tests/beamformer/
beamformer_sum_test.py:64-77).  They differ in which beams they emit, header edits and ring sizing only.

The reference relies on the Beamform block having initialised bifrost's library-global beamformer state
(beamform_sum_beams_block.py:186-187).  libxeng keeps no global state: each block owns a handle sized from the
sequence header.
"""
import json
import time

from .. import ops
from .. import runtime as rt
from .block_base import Block


class BeamPowerBlock(Block):
    """Subclasses provide ``_plan(ihdr)`` -> dict(nchan, nbeam_in, igulp_size, ogulp_size, ohdr, ring_size,
    beam_id or None)."""

    def _init_common(self, ntime_gulp, ntime_sum, gpu):
        self.ntime_gulp, self.ntime_sum, self.gpu = ntime_gulp, ntime_sum, gpu
        assert ntime_gulp % ntime_sum == 0
        self.ntime_blocks = ntime_gulp // ntime_sum
        self._handle = None
        self._handle_dims = None
        rt.set_device(self.gpu)

    def _engine(self, nchan, nbeam):
        if self._handle_dims != (nchan, nbeam):
            if self._handle is not None:
                self._handle.close()
            self._handle = ops.Beamformer(gpu=self.gpu, ninput=1, nchan=nchan, ntime=self.ntime_gulp, nbeam=nbeam)
            self._handle_dims = (nchan, nbeam)
        return self._handle

    def main(self):
        rt.set_core(self.core)
        rt.set_device(self.gpu)
        self.bind_proclog.update({'ncore': 1, 'core0': rt.get_core(), 'ngpu': 1, 'gpu0': rt.get_device()})
        with self.oring.begin_writing() as oring:
            for iseq in self.iring.read(guarantee=self.guarantee):
                ihdr = json.loads(iseq.header.tostring())
                self.sequence_proclog.update(ihdr)
                plan = self._plan(ihdr)
                igulp_size, ogulp_size = plan['igulp_size'], plan['ogulp_size']
                engine = self._engine(plan['nchan'], plan['nbeam_in'])
                self.oring.resize(plan['ring_size'])
                prev = time.time()
                with oring.begin_sequence(time_tag=iseq.time_tag, header=json.dumps(plan['ohdr'])) as oseq:
                    for ispan in iseq.read(igulp_size):
                        if ispan.size < igulp_size:
                            continue
                        t = time.time()
                        acquire_time, prev = t - prev, t
                        with oseq.reserve(ogulp_size) as ospan:
                            t = time.time()
                            reserve_time, prev = t - prev, t
                            # the integrated powers are written straight into the output span (device memory, or
                            # pinned host memory the GPU can address), so there is no second copy
                            if plan['beam_id'] is None:
                                engine.integrate(ispan.data, ospan.data, self.ntime_sum)
                            else:
                                engine.integrate_single(ispan.data, ospan.data, self.ntime_sum, plan['beam_id'])
                            engine.sync()
                        t = time.time()
                        process_time, prev = t - prev, t
                        self.perf_proclog.update({'acquire_time': acquire_time, 'reserve_time': reserve_time,
                                                  'process_time': process_time,
                                                  'gbps': 8 * igulp_size / process_time / 1e9})
