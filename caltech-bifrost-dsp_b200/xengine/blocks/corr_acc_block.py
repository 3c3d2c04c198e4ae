"""CorrAcc: long-term accumulation of Corr's integrations on the device.

Drop-in for ``lwa352_pipeline.blocks.corr_acc_block.CorrAcc`` (reference: pipeline/lwa352_pipeline/blocks/
corr_acc_block.py:17-336): same constructor signature, command keys, header keys (``upstream_acc_len``,
``acc_len``, ``seq0``) and output layout (the input layout, int32 [2][nchan][words_per_chan]).  The reference's
runtime-compiled ``BFMap("a = b")`` / ``BFMap("a += b")`` (corr_acc_block.py:304,306) become ``xeng_acc``.
"""
import json
import time

from .. import ops
from .. import runtime as rt
from ._integration import IntegrationClock
from .block_base import Block


class CorrAcc(Block):
    def __init__(self, log, iring, oring, guarantee=True, core=-1, nchan=192, npol=2, nstand=352, acc_len=24000,
                 gpu=-1, etcd_client=None, autostartat=0):
        super(CorrAcc, self).__init__(log, iring, oring, guarantee, core, etcd_client=etcd_client)
        self.nchan, self.npol, self.nstand = nchan, npol, nstand
        self.matlen = nchan * (nstand // 2 + 1) * (nstand // 4) * npol * npol * 4
        self.gpu = gpu
        rt.set_device(self.gpu)
        self.igulp_size = self.matlen * 8
        self.ogulp_size = self.igulp_size
        self.accdata = rt.empty(self.igulp_size // 4, 'i32', 'cuda')      # the integration buffer
        self.define_command_key('start_time', type=int, initial_val=autostartat)
        self.define_command_key('acc_len', type=int, initial_val=acc_len)

    def _check_alignment(self, clock, upstream_acc_len, upstream_start):
        if clock.acc_len % upstream_acc_len != 0:
            self.log.error("CORRACC >> Requested acc_len %d incompatible with upstream integration %d"
                           % (clock.acc_len, upstream_acc_len))
        if clock.acc_len != 0 and ((clock.start_time - upstream_start) % upstream_acc_len != 0):
            self.log.error("CORRACC >> Requested start_time %d incompatible with upstream integration %d"
                           % (clock.start_time, upstream_acc_len))

    def main(self):
        rt.set_core(self.core)
        rt.set_device(self.gpu)
        self.bind_proclog.update({'ncore': 1, 'core0': rt.get_core()})
        self.oring.resize(self.ogulp_size)
        clock = IntegrationClock(recover_skip=2, asap="now")
        nwords = self.igulp_size // 4
        oseq = None
        time_tag = 1
        process_time = acquire_time = reserve_time = 0.
        self.update_stats({'state': 'starting'})
        with self.oring.begin_writing() as oring:
            prev = time.time()
            self.update_pending = True
            for iseq in self.iring.read(guarantee=self.guarantee):
                ihdr = json.loads(iseq.header.tostring())
                ohdr = dict(ihdr)
                now = upstream_start = ihdr['seq0']
                step = ihdr['acc_len']                       # one input span = one upstream integration
                ohdr['upstream_acc_len'] = step
                self.sequence_proclog.update(ohdr)
                if clock.resequence(now):
                    self.log.info("CORRACC >> Recovering start time set to %d. Accumulating %d samples"
                                  % (clock.start_time, clock.acc_len))
                    self._check_alignment(clock, step, upstream_start)
                    ohdr['acc_len'], ohdr['seq0'] = clock.acc_len, clock.start_time
                for ispan in iseq.read(self.igulp_size):
                    if ispan.size < self.igulp_size:
                        continue
                    if self.update_pending:
                        self.update_command_vals()
                        clock.retarget(self.command_vals['start_time'], self.command_vals['acc_len'], now)
                        self.log.info("CORRACC >> New start time at %d. Accumulation: %d samples"
                                      % (clock.start_time, clock.acc_len))
                        self._check_alignment(clock, step, upstream_start)
                        ohdr['acc_len'], ohdr['seq0'] = clock.acc_len, clock.start_time
                    self.stats.update({'curr_sample': now})
                    self.update_stats()
                    if clock.acc_len == 0:                   # stop condition
                        self.update_stats({'state': 'stopped'})
                        if oseq:
                            oseq.end()
                        oseq = None
                        clock.stop()
                        now += step
                        continue
                    if clock.begins(now, step):
                        if oseq:
                            oseq.end()
                        self.sequence_proclog.update(ohdr)
                        oseq = oring.begin_sequence(time_tag=time_tag, header=json.dumps(ohdr),
                                                    nringlet=iseq.nringlet)
                        time_tag += 1
                        self.log.info("CORRACC >> Start time %d reached. Accumulating to %d (upstream accumulation: %d)"
                                      % (clock.start_time, clock.last, step))
                    if not clock.running:
                        self.update_stats({'state': 'waiting_start_missed' if now > clock.start_time else 'waiting'})
                        now += step
                        continue
                    self.update_stats({'state': 'running'})
                    t = time.time()
                    acquire_time, prev = t - prev, t
                    # acc = in on the first span of an integration, acc += in afterwards
                    ops.acc(self.accdata, ispan.data, nwords, first=(now == clock.first))
                    rt.stream_synchronize()                  # the input span may be released after this iteration
                    t = time.time()
                    process_time += t - prev
                    prev = t
                    if now == clock.last:
                        self.log.debug("CORRACC > Last accumulation input")
                        ospan = rt.WriteSpan(oseq.ring, self.ogulp_size, nonblocking=False)
                        t = time.time()
                        reserve_time, prev = t - prev, t
                        rt.copy_array(ospan.data_view('i32'), self.accdata)
                        rt.stream_synchronize()              # the copy must land before the span is committed
                        ospan.close()
                        t = time.time()
                        process_time += t - prev
                        prev = t
                        self.perf_proclog.update({'acquire_time': acquire_time, 'reserve_time': reserve_time,
                                                  'process_time': process_time})
                        self.update_stats({'last_end_sample': now})
                        process_time = 0.
                        clock.advance(step)
                    now += step
            if oseq:
                oseq.end()
