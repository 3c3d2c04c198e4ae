"""Corr: cross-correlate 4+4-bit voltages on the B200 tensor cores, output in xGPU's triangular order.

Drop-in for ``lwa352_pipeline.blocks.corr_block.Corr`` (reference: pipeline/lwa352_pipeline/blocks/
corr_block.py:60-472): same constructor signature, command keys (``start_time``, ``acc_len``), ring header
keys and output byte layout.  Where the reference calls ``_bf.bfXgpuInitialize`` / ``bfXgpuKernel`` /
``bfXgpuGetOrder`` (corr_block.py:253,445,331) this block calls libxeng through ``xengine.ops``.

Differences that do not change the data products:
  * the integration is accumulated straight into the reserved output span (``xeng_corr_accumulate``), so there
    is no library-internal accumulator to go stale when an input sequence ends mid-integration;
  * ``acc_len = 0`` stops cleanly (the reference falls through to ``WriteSpan(None.ring, ...)``,
    corr_block.py:421-435);
  * ``test=True`` cross-checks every dump on the GPU against this library's second, CUDA-core implementation. That is
    a fast on-line consistency check between two kernels by the same author; it is NOT the equivalent of the
    reference's in-line numpy ``_test`` / ``_compare`` (corr_block.py:265-315, ~10 minutes per gulp), whose role --
    parity with the reference's formulation -- is played by the oracle in ``tests/`` (restated as
    ``oracle.corr_loop_literal`` and pinned to the reference's own ``Corr._test`` output).

Optional keyword arguments at the END of the signature (all default to the reference's behaviour):

``fast_oring=None, nchan_sum=4, fast_nvis=4704`` -- fused fast dump (BASELINE configs[4]).  Every integration is also
written to ``fast_oring`` as CorrSubsel would write it (int32 [nchan/nchan_sum][nvis][2], CorrSubsel's header keys,
corr_subsel_block.py:272-285) -- gathered inside the correlator kernel's epilogue (``xeng_corr_accumulate_subsel``), so
no CorrSubsel block and no second pass over the 383 MB matrix are needed.  The selection is set with
``update_fast_subsel(baselines)`` (same [[stand0, pol0], [stand1, pol1]] lists as CorrSubsel's ``baselines`` key) and
takes effect at the next integration boundary.  The ring may live in any space: for anything but 'cuda' (production
uses a 'cuda_host' ring, lwa352-pipeline.py:155) the kernel gathers into a device scratch buffer that is copied into
the span when the integration closes -- the kernel's reductions never cross PCIe.

``acc_oring=None, slow_acc_len=None`` -- fused CorrAcc.  With ``acc_oring`` set (and ``oring=None``), the long
integration CorrAcc would form from this block's dumps (lwa352-pipeline.py:232-276: Corr -> CorrAcc) is accumulated
by the SAME launch straight into a span of ``acc_oring``: ``slow_acc_len`` samples per dump (a multiple of
``acc_len``), CorrAcc's header keys (``upstream_acc_len``, ``acc_len``, ``seq0``; corr_acc_block.py:216,255-256).
``acc_len`` then only sets the fast dump's cadence.  One launch per gulp does: long accumulator read-modify-write +
fast-dump gather -- the path bench.py's headline number times.

``gulps_per_span=1`` -- read ``gulps_per_span`` gulps per input span and launch them back to back, waiting for the
device once per span (and at integration ends) instead of once per gulp: the host bookkeeping of one gulp then
overlaps the kernel of the previous one.
"""
import contextlib
import json
import time

import numpy as np

from .. import ops
from .. import runtime as rt
from ._integration import IntegrationClock
from .block_base import Block


def tri_index(i, j):
    """Index into a lower-triangular matrix stored row by row (corr_block.py:27-35)."""
    return (i * (i + 1)) // 2 + j


def regtile_index(in0, in1, nstand):
    """Word index of the product of inputs ``in0`` and ``in1`` (stand of in1 >= stand of in0) inside one channel of
    the xGPU register-tile-order buffer (corr_block.py:37-58)."""
    a0, a1, p0, p1 = in0 >> 1, in1 >> 1, in0 & 1, in1 & 1
    quadrant = 2 * (a0 & 1) + (a1 & 1)
    cells_per_quadrant = (nstand // 2 + 1) * nstand // 4
    cell = quadrant * cells_per_quadrant + tri_index(a1 // 2, a0 // 2)
    return 4 * cell + 2 * p1 + p0


class Corr(Block):
    def __init__(self, log, iring, oring, ntime_gulp=2500, guarantee=True, core=-1, nchan=192, npol=2, nstand=352,
                 acc_len=2400, gpu=-1, test=False, etcd_client=None, autostartat=0, ant_to_input=None,
                 fast_oring=None, nchan_sum=4, fast_nvis=4704, acc_oring=None, slow_acc_len=None, gulps_per_span=1):
        assert (acc_len % ntime_gulp == 0), "Acculmulation length must be a multiple of gulp size"
        if acc_oring is not None:
            if oring is not None:
                raise ValueError("Corr: with acc_oring (fused CorrAcc) the per-integration matrix is not produced; pass oring=None")
            if not slow_acc_len or slow_acc_len % acc_len:
                raise ValueError("Corr: slow_acc_len must be a positive multiple of acc_len")
        elif oring is None:
            raise ValueError("Corr: need oring or acc_oring")
        super(Corr, self).__init__(log, iring, oring, guarantee, core, etcd_client=etcd_client)
        self.ntime_gulp, self.nchan, self.npol, self.nstand = ntime_gulp, nchan, npol, nstand
        self.matlen = nchan * (nstand // 2 + 1) * (nstand // 4) * npol * npol * 4
        self.gpu = gpu
        self.test = test
        rt.set_device(self.gpu)

        self.size_proclog.update({'nseq_per_gulp': self.ntime_gulp})
        self.igulp_size = self.ntime_gulp * nchan * nstand * npol      # one byte per 4+4-bit sample
        self.ogulp_size = self.matlen * 8                              # int32 real plane + int32 imaginary plane

        self.define_command_key('start_time', type=int, initial_val=autostartat,
                                condition=lambda x: (x == -1) or (x % self.ntime_gulp == 0))
        self.define_command_key('acc_len', type=int, initial_val=acc_len,
                                condition=lambda x: x % self.ntime_gulp == 0)
        self.update_stats({'xgpu_acc_len': self.ntime_gulp})

        try:
            self._xcorr = ops.XCorr(nstand=nstand, npol=npol, nchan=nchan, ntime_gulp=ntime_gulp, gpu=gpu)
        except Exception as e:
            self.log.error("xeng_corr_create failed: %s" % e)
            raise RuntimeError(str(e))
        assert self._xcorr.matlen == self.matlen
        self._xcheck = None     # CUDA-core cross-check engine, created on first use when test=True

        self.antpol_to_input = np.zeros([nstand, npol], dtype=np.int32)
        self.antpol_to_bl = np.zeros([nstand, nstand, npol, npol], dtype=np.int32)
        self.bl_is_conj = np.zeros([nstand, nstand, npol, npol], dtype=np.int32)
        if ant_to_input is not None:
            self.update_baseline_indices(ant_to_input)

        # fused fast-dump path
        self.fast_oring, self.nchan_sum, self.fast_nvis = fast_oring, nchan_sum, fast_nvis
        self.fast_gulp_size = (nchan // nchan_sum) * fast_nvis * 8
        self._fast_baselines = [[[i % nstand, 0], [i % nstand, 0]] for i in range(fast_nvis)]   # corr_subsel_block.py:230
        self._fast_pending = fast_oring is not None
        self._fast_dev = None        # device scratch of the fast dump when fast_oring is not a 'cuda' ring
        if fast_oring is not None and getattr(fast_oring, 'space', 'cuda') != 'cuda':
            self._fast_dev = rt.zeros(self.fast_gulp_size // 4, 'i32', 'cuda')
        self.acc_oring, self.slow_acc_len = acc_oring, slow_acc_len
        self.gulps_per_span = max(1, int(gulps_per_span))

    def update_fast_subsel(self, baselines):
        """Selection of the fused fast dump: ``fast_nvis`` entries [[stand0, pol0], [stand1, pol1]]; loaded at the
        next integration boundary (a new output sequence on the fast ring carries it in its header)."""
        if len(baselines) != self.fast_nvis:
            raise ValueError("expected %d baselines, got %d" % (self.fast_nvis, len(baselines)))
        with self._control_lock:
            self._fast_baselines = baselines
            self._fast_pending = True

    def _load_fast_selection(self):
        with self._control_lock:
            baselines = self._fast_baselines
            self._fast_pending = False
        bl = np.asarray(baselines, dtype=np.int64).reshape(self.fast_nvis, 2, 2)
        s0, p0, s1, p1 = bl[:, 0, 0], bl[:, 0, 1], bl[:, 1, 0], bl[:, 1, 1]
        self._xcorr.set_subsel(self.antpol_to_bl[s0, s1, p0, p1], self.bl_is_conj[s0, s1, p0, p1], self.nchan_sum)
        return baselines

    def _fast_header(self, ohdr, baselines):
        """CorrSubsel's output header (corr_subsel_block.py:272-285) for the sequence described by ``ohdr``."""
        fhdr = dict(ohdr)
        fhdr['nchan'] = ohdr['nchan'] // self.nchan_sum
        fhdr['nvis'] = self.fast_nvis
        fhdr['sfreq'] = (ohdr['sfreq'] + (self.nchan_sum - 1) * (ohdr['bw_hz'] / ohdr['nchan'])) / self.nchan_sum
        fhdr['baselines'] = baselines
        fhdr['nchan_sum'] = self.nchan_sum
        return fhdr

    def update_baseline_indices(self, ant_to_input):
        """Fill ``antpol_to_bl`` / ``bl_is_conj`` ([nstand, nstand, npol, npol]) from a [nstand, npol] map of
        correlator input numbers: the word index inside one channel of the output buffer that holds the product
        of (stand0, pol0) and (stand1, pol1), and whether it must be conjugated to read as
        ``v(s0,p0) conj(v(s1,p1))``.  In place, so downstream blocks constructed with these arrays see updates."""
        rt.set_core(self.core)
        self.antpol_to_input[...] = ant_to_input
        a2bl, conj = ops.get_order(self.antpol_to_input)
        self.antpol_to_bl[...] = a2bl
        self.bl_is_conj[...] = conj

    # ------------------------------------------------------------------ test mode
    def _crosscheck_gulp(self, idata, first):
        """test=True: repeat the gulp on the CUDA-core kernel into a private accumulator."""
        if self._xcheck is None:
            self._xcheck = ops.XCorr(nstand=self.nstand, npol=self.npol, nchan=self.nchan,
                                     ntime_gulp=self.ntime_gulp, gpu=self.gpu)
            self._xcheck.set_impl(1)
            self._xcheck_acc = rt.zeros(2 * self.matlen, 'i32', 'cuda')
        self._xcheck.accumulate(idata, self._xcheck_acc, first)

    def _crosscheck_compare(self, odata):
        import torch
        self._xcheck.sync()
        got = odata.view(torch.int32) if isinstance(odata, torch.Tensor) else torch.as_tensor(odata).view(torch.int32)
        ok = bool(torch.equal(got.reshape(-1), self._xcheck_acc.reshape(-1)))
        self.update_stats({'test_match': ok})
        (self.log.info if ok else self.log.error)("CORR >> tensor-core vs CUDA-core dump match: %s" % ok)
        return ok

    # ------------------------------------------------------------------ main loop
    def _set_state(self, state):
        if self.stats.get('state') != state:        # published on change only: a ProcLog update per gulp is pure overhead
            self.update_stats({'state': state})

    def main(self):
        rt.set_core(self.core)
        rt.set_device(self.gpu)
        self.bind_proclog.update({'ncore': 1, 'core0': rt.get_core()})
        fused_acc = self.acc_oring is not None
        main_ring = self.acc_oring if fused_acc else self.oring
        main_ring.resize(self.ogulp_size)
        step = self.ntime_gulp
        clock = IntegrationClock(recover_skip=10, asap="boundary")
        time_tag = 1
        self.update_stats({'state': 'starting'})
        with contextlib.ExitStack() as stack:
            oring = stack.enter_context(main_ring.begin_writing())
            foring = None
            if self.fast_oring is not None:
                self.fast_oring.resize(self.fast_gulp_size)
                foring = stack.enter_context(self.fast_oring.begin_writing())
            prev = time.time()
            self.update_pending = True          # load the constructor's start_time / acc_len on the first gulp
            oseq = ospan = None
            foseq = fspan = None
            slow_first = slow_last = None       # fused CorrAcc: window of the long integration
            for iseq in self.iring.read(guarantee=self.guarantee):
                self.log.info('CORR >> new input sequence!')
                ihdr = json.loads(iseq.header.tostring())
                now = ihdr['seq0']
                ohdr = dict(ihdr)
                # the output is xGPU-ordered, not [stand, pol]: the input maps do not describe it
                ohdr.pop('ant_to_input', None)
                ohdr.pop('input_to_ant', None)
                if clock.resequence(now):
                    self.log.info("CORR >> Recovering start time set to %d. Accumulating %d samples"
                                  % (clock.start_time, clock.acc_len))
                    ohdr['acc_len'], ohdr['seq0'] = clock.acc_len, clock.start_time
                self.sequence_proclog.update(ohdr)
                ospan = fspan = None            # an integration cut short by the end of a sequence is dropped
                process_time = acquire_time = reserve_time = 0.
                for ispan in iseq.read(self.gulps_per_span * self.igulp_size):
                    ngulp = ispan.size // self.igulp_size
                    if ngulp == 0:
                        self.log.info("CORR >>> Ignoring final gulp (expected %d bytes but got %d)"
                                      % (self.igulp_size, ispan.size))
                        continue
                    in_flight = False           # launches of this span the device may still be running
                    closing, last_ends = [], []  # output spans to commit / integrations ended, once the device is done
                    for k in range(ngulp):
                        idata = ispan.data if ngulp == 1 and ispan.size == self.igulp_size else \
                            ispan.data[k * self.igulp_size:(k + 1) * self.igulp_size]
                        if self.update_pending:
                            self.update_command_vals()
                            clock.retarget(self.command_vals['start_time'], self.command_vals['acc_len'], now)
                            self.log.info("CORR >> New start time set to %d. Accumulating %d samples"
                                          % (clock.start_time, clock.acc_len))
                            ohdr['acc_len'], ohdr['seq0'] = clock.acc_len, clock.start_time
                            if fused_acc and clock.acc_len and self.slow_acc_len % clock.acc_len:
                                self.log.error("CORR >> acc_len %d does not divide slow_acc_len %d"
                                               % (clock.acc_len, self.slow_acc_len))
                        self.stats['curr_sample'] = now         # published with the next update_stats
                        if clock.acc_len == 0:      # stop condition
                            self._set_state('stopped')
                            if oseq:
                                oseq.end()
                            if foseq:
                                foseq.end()
                            oseq = ospan = foseq = fspan = None
                            clock.stop()
                            now += step
                            continue
                        if clock.begins(now, step):
                            self.log.info("CORR >> Start time %d reached." % clock.start_time)
                            if oseq:
                                oseq.end()
                            mhdr = dict(ohdr)
                            if fused_acc:           # CorrAcc's output header (corr_acc_block.py:216,255-256)
                                mhdr.update({'upstream_acc_len': clock.acc_len, 'acc_len': self.slow_acc_len, 'seq0': now})
                                slow_first, slow_last = now, now + self.slow_acc_len - step
                                ospan = None
                            self.sequence_proclog.update(mhdr)
                            oseq = oring.begin_sequence(time_tag=time_tag, header=json.dumps(mhdr),
                                                        nringlet=iseq.nringlet)
                            if foring is not None:
                                if foseq:
                                    foseq.end()
                                fhdr = self._fast_header(ohdr, self._load_fast_selection())
                                foseq = foring.begin_sequence(time_tag=time_tag, header=json.dumps(fhdr),
                                                              nringlet=iseq.nringlet)
                            time_tag += 1
                        if not clock.running:
                            self._set_state('waiting')
                            now += step
                            continue
                        self._set_state('running')
                        t = time.time()
                        acquire_time, prev = t - prev, t
                        is_first, is_last = (now == clock.first), (now == clock.last)
                        acc_first, acc_last = (now == slow_first, now == slow_last) if fused_acc else (is_first, is_last)
                        if acc_first:
                            ospan = rt.WriteSpan(oseq.ring, self.ogulp_size, nonblocking=False)
                        if is_first and foring is not None:
                            if self._fast_pending:      # new selection: new fast sequence with an updated header
                                foseq.end()
                                fhdr = self._fast_header(ohdr, self._load_fast_selection())
                                fhdr['seq0'] = now
                                foseq = foring.begin_sequence(time_tag=time_tag, header=json.dumps(fhdr),
                                                              nringlet=iseq.nringlet)
                                time_tag += 1
                            fspan = rt.WriteSpan(foseq.ring, self.fast_gulp_size, nonblocking=False)
                        if acc_first or is_first:
                            t = time.time()
                            reserve_time, prev = t - prev, t
                        if ospan is None:
                            self.log.error("CORR: trying to write to not-yet-opened ospan")
                            now += step
                            continue
                        # device boundary (replaces _bf.bfXgpuKernel, corr_block.py:445): ospan (+)= X(ispan)
                        if fspan is not None:       # the same launch also gathers the fast-dump selection
                            ftarget = self._fast_dev if self._fast_dev is not None else fspan.data
                            self._xcorr.accumulate_subsel(idata, ospan.data, acc_first, ftarget, is_first)
                        else:
                            self._xcorr.accumulate(idata, ospan.data, acc_first)
                        if self.test:
                            self._crosscheck_gulp(idata, acc_first)
                        in_flight = True
                        if is_last and fspan is not None and self._fast_dev is not None:
                            # non-device fast ring: the kernel gathered into device scratch; the copy into the span is
                            # queued behind it (and ahead of the next gulp's launch, which zeroes the scratch again)
                            rt.copy_array(fspan.data, self._fast_dev)
                        if acc_last:
                            closing.append((ospan, True))
                            ospan = None
                            if fused_acc:
                                slow_first, slow_last = slow_last + step, slow_last + self.slow_acc_len
                        if is_last:
                            if fspan is not None:
                                closing.append((fspan, False))
                                fspan = None
                            last_ends.append(now)
                            clock.advance(step)
                        now += step
                    # ---- end of the input span: ONE wait for the device covers every launch of the span; the output
                    # spans whose integrations ended inside it are committed afterwards, in order (a span is committed
                    # only once the device has written it; the input span is released when this iteration ends)
                    if in_flight:
                        self._xcorr.sync()
                        t = time.time()
                        process_time += t - prev
                        prev = t
                    for sp, is_matrix in closing:
                        if is_matrix and self.test:
                            self._crosscheck_compare(sp.data)
                        sp.close()
                    del closing[:]
                    if last_ends:
                        nint = len(last_ends)
                        gbps = 8 * nint * clock.acc_len * ihdr['nchan'] * ihdr['nstand'] * ihdr['npol'] / max(process_time, 1e-9) / 1e9
                        self.perf_proclog.update({'acquire_time': acquire_time, 'reserve_time': reserve_time,
                                                  'process_time': process_time / nint, 'gbps': gbps})
                        self.update_stats({'last_end_sample': last_ends[-1], 'throughput': gbps})
                        process_time = 0.
                        del last_ends[:]
                if oseq:
                    oseq.end()
                if foseq:
                    foseq.end()
                oseq = foseq = None
