"""Copy: move gulps from one ring to another -- in production the H2D stage in front of the X-engine
(capture ring in pinned host memory -> ``gpu-input`` ring, lwa352-pipeline.py:205-214).

Drop-in for ``lwa352_pipeline.blocks.copy_block.Copy`` (reference: pipeline/lwa352_pipeline/blocks/
copy_block.py:17-160): same constructor signature, headers copied unchanged, partial final gulps ignored, output ring
sized ``buffer_multiplier`` gulps per span.

The reference copies one gulp per loop iteration with bifrost's ``copy_array`` (copy_block.py:143).  Here the block reads
``buffer_multiplier`` gulps per input span, enqueues one asynchronous copy per gulp on a private stream, and commits
each output gulp as soon as ITS copy has landed: the link keeps streaming the rest of the span while downstream
blocks already work on the first gulp, and the only time the copy engine idles is the host's turn-around between two
input spans.  The input span is held until its last copy is done (a ring may reuse a released span at once).
"""
import time

from .. import runtime as rt
from .block_base import Block


class Copy(Block):
    def __init__(self, log, iring, oring, ntime_gulp=2500, buffer_multiplier=1, guarantee=True, core=-1,
                 nbyte_per_time=184 * 352 * 2, gpu=-1, buf_size_gbytes=None):
        super(Copy, self).__init__(log, iring, oring, guarantee, core, etcd_client=None)
        rt.set_core(self.core)
        self.ntime_gulp, self.gpu = ntime_gulp, gpu
        rt.set_device(self.gpu)
        self.buffer_multiplier = max(1, int(buffer_multiplier))
        self.size_proclog.update({'nseq_per_gulp': self.ntime_gulp})
        self.igulp_size = self.ntime_gulp * nbyte_per_time
        unit = self.igulp_size * self.buffer_multiplier
        if buf_size_gbytes is None:                       # copy_block.py:110-113
            self.buf_size = 4 * unit
        else:
            self.buf_size = int(1e9 * buf_size_gbytes) // unit * unit
        self.oring.resize(unit, total_span=self.buf_size)

    def main(self):
        rt.set_core(self.core)
        rt.set_device(self.gpu)
        self.bind_proclog.update({'ncore': 1, 'core0': rt.get_core()})
        lane = rt.CopyLane()
        g = self.igulp_size
        with self.oring.begin_writing() as oring:
            for iseq in self.iring.read(guarantee=self.guarantee):
                prev = time.time()
                bytes_copied, acquire_time, reserve_time, process_time = 0, 0., 0., 0.
                with oring.begin_sequence(time_tag=iseq.time_tag, header=iseq.header.tostring().decode("utf-8"),
                                          nringlet=iseq.nringlet) as oseq:
                    for ispan in iseq.read(self.buffer_multiplier * g):
                        ngulp = ispan.size // g
                        if ngulp == 0:
                            continue                      # ignore a partial final gulp (copy_block.py:134-135)
                        t = time.time()
                        acquire_time, prev = acquire_time + t - prev, t
                        pending = []
                        for k in range(ngulp):
                            ospan = rt.WriteSpan(oseq.ring, g, nonblocking=False)
                            src = ispan.data if ngulp == 1 and ispan.size == g else ispan.data[k * g:(k + 1) * g]
                            pending.append((ospan, lane.copy(ospan.data, src)))
                        t = time.time()
                        reserve_time, prev = reserve_time + t - prev, t
                        for ospan, ticket in pending:     # commit in order, each as soon as its copy has landed
                            lane.wait(ticket)
                            ospan.close()
                        t = time.time()
                        process_time, prev = process_time + t - prev, t
                        bytes_copied += ngulp * g
                        if bytes_copied > 10e9:
                            self.perf_proclog.update({'acquire_time': acquire_time, 'reserve_time': reserve_time,
                                                      'process_time': process_time,
                                                      'gbps': 8 * bytes_copied / max(process_time, 1e-9) / 1e9})
                            bytes_copied, acquire_time, reserve_time, process_time = 0, 0., 0., 0.
                if bytes_copied:
                    self.perf_proclog.update({'acquire_time': acquire_time, 'reserve_time': reserve_time,
                                              'process_time': process_time,
                                              'gbps': 8 * bytes_copied / max(process_time, 1e-9) / 1e9})
