"""Beamform: voltage beamformer, out[chan, beam, t] = sum_i w[chan, beam, i] * v[t, chan, i].

Drop-in for ``lwa352_pipeline.blocks.beamform_block.Beamform`` (reference: pipeline/lwa352_pipeline/blocks/
beamform_block.py:18-461): same constructor signature, the single ``coeffs`` command key carrying ``calgains`` /
``beamcoeffs`` dictionaries (beamform_block.py:320-362), same output header (``nstand = nbeam``, ``nbit = 32``,
``npol = 1``, ``complex = True``, ``nbeam``) and output layout cf32 [nchan][nbeam][ntime].
``_bf.bfBeamformInitialize`` / ``bfBeamformRun`` (beamform_block.py:251-253,449) become ``xeng_bf_create`` /
``xeng_bf_run``.  With ``ntime_sum`` set (the reference's "experimental" mode) the block emits integrated dual-pol
powers f32 [nbeam/2][ntime_gulp/ntime_sum][nchan][4] instead of voltages: same byte count per gulp as the
reference reserves (beamform_block.py:373).
"""
import json
import time

import numpy as np

from .. import ops
from .. import runtime as rt
from .block_base import COMMAND_INVALID, COMMAND_OK, Block


class Beamform(Block):
    def __init__(self, log, iring, oring, nchan=256, nbeam=1, ninput=352 * 2, ntime_gulp=2500, ntime_sum=None,
                 guarantee=True, core=-1, gpu=-1, etcd_client=None):
        super(Beamform, self).__init__(log, iring, oring, guarantee, core, etcd_client=etcd_client)
        self.ntime_gulp, self.gpu, self.ntime_sum = ntime_gulp, gpu, ntime_sum
        if ntime_sum is not None:
            assert ntime_gulp % ntime_sum == 0
            self.ntime_blocks = ntime_gulp // ntime_sum
        else:
            self.ntime_blocks = ntime_gulp
        self.nchan, self.nbeam, self.ninput = nchan, nbeam, ninput
        self.freqs = np.zeros(self.nchan, dtype=np.float32)
        rt.set_device(self.gpu)

        shape = (nchan, nbeam, ninput)
        self.cal_gains = np.ones(shape, dtype=np.complex64)       #: per-input calibration gains
        self.gains_cpu_new = np.zeros(shape, dtype=np.complex64)  #: latest coefficients, waiting for their load time
        self.gains_cpu = np.zeros(shape, dtype=np.complex64)      #: coefficients to be copied to the device
        self.gains_gpu = rt.zeros(shape, np.complex64, 'cuda')    #: coefficients the kernel reads
        self.gains_load_sample = np.zeros(nbeam)                  #: sample at which gains_cpu_new[b] becomes active

        self.define_command_key('coeffs', type=dict, initial_val={})
        for b in range(self.nbeam):
            self.update_stats({'cal_gains%d' % b: [False] * ninput})

        if ntime_sum is not None:
            self.log.warning("Running Beamform block with ntime_sum != None is experimental!")
        self._bf = ops.Beamformer(gpu=self.gpu, ninput=ninput, nchan=nchan, ntime=ntime_gulp, nbeam=nbeam,
                                  ntime_blocks=self.ntime_blocks if ntime_sum is not None else 0)

    # ------------------------------------------------------------------ control
    def _etcd_callback(self, watchresponse):
        """Coefficient commands arrive often and all share one key, so each is applied (update_command_vals) as
        soon as it is received instead of waiting for ``main()`` (beamform_block.py:267-317)."""
        rt.set_core(self.core)
        with self._control_lock:
            for event in watchresponse.events:
                try:
                    seq_id, payload = self._decode_command(event.value)
                except ValueError:
                    self.log.exception("BEAMFORM >> Failed to JSON-decode event %s" % str(event.value))
                    self._send_command_response("0", False, "JSON-decode failed!")
                    continue
                if not isinstance(payload, dict):
                    self._send_command_response(seq_id, False, payload)
                    continue
                try:
                    rc = self._process_commands(payload, set_pending_flag=False)
                except Exception:
                    rc = COMMAND_INVALID
                self.update_stats({'last_cmd_response': rc})
                self.update_command_vals()
                self._send_command_response(seq_id, rc == COMMAND_OK, str(rc))

    def update_command_vals(self):
        """Apply the pending ``coeffs`` command (called with the control lock held):

        ``calgains``   {'type','input_id','beam_id','data': [re0, im0, re1, im1, ...] per channel}
        ``beamcoeffs`` {'type','beam_id','data': {'delays': [ns per input], 'amps': [per input]}, 'load_sample'}
                       -> gains[:, b, :] = amps * exp(2 pi i freqs delays 1e-9) * cal_gains[:, b, :]
        """
        self.command_vals.update(self._pending_command_vals)
        cal_changed = False
        for v in self._pending_command_vals.values():
            try:
                kind = v['type']
                if kind == 'calgains':
                    i, b = v['input_id'], v['beam_id']
                    data = np.array(v['data'])
                    self.cal_gains[:, b, i] = data[0::2] + 1j * data[1::2]
                    self.stats['cal_gains%d' % b][i] = True
                    cal_changed = True
                elif kind == 'beamcoeffs':
                    b = v['beam_id']
                    delays_ns = np.array(v['data']['delays'])
                    amps = np.array(v['data']['amps'])
                    phases = np.exp(1j * 2 * np.pi * self.freqs[:, None] * delays_ns * 1e-9)
                    self.gains_cpu_new[:, b, :] = amps * phases * self.cal_gains[:, b, :]
                    self.gains_load_sample[b] = v.get('load_sample', -1)    # default: load immediately
                    self.update_pending = True      # only beam coefficients trigger a device update
            except (KeyError, TypeError):
                self.log.error("BEAMFORM >> Failed to parse command")
        self.update_stats(self.command_vals)
        if cal_changed:
            self.update_stats({'cal_gains%d' % b: self.stats['cal_gains%d' % b] for b in range(self.nbeam)})

    # ------------------------------------------------------------------ main loop
    def main(self):
        rt.set_core(self.core)
        rt.set_device(self.gpu)
        self.bind_proclog.update({'ncore': 1, 'core0': rt.get_core(), 'ngpu': 1, 'gpu0': rt.get_device()})
        igulp_size = self.ntime_gulp * self.nchan * self.ninput            # 4+4-bit samples
        ogulp_size = self.ntime_blocks * self.nchan * self.nbeam * 8       # cf32 voltages (or 4 x f32 powers)
        self.oring.resize(ogulp_size)
        with self.oring.begin_writing() as oring:
            for iseq in self.iring.read(guarantee=self.guarantee):
                self.update_pending = True      # frequencies may have changed: re-examine the coefficients
                copy_pending = True
                ihdr = json.loads(iseq.header.tostring())
                self.sequence_proclog.update(ihdr)
                now = ihdr['seq0']
                nchan, sfreq = ihdr['nchan'], ihdr['sfreq']
                chan_bw = ihdr['bw_hz'] / nchan
                assert nchan == self.nchan
                assert self.ninput == ihdr['nstand'] * ihdr['npol']
                self.freqs = sfreq + chan_bw * np.arange(nchan)
                ohdr = dict(ihdr)
                ohdr.update({'nstand': self.nbeam, 'nbit': 32, 'npol': 1, 'complex': True, 'nbeam': self.nbeam})
                prev = time.time()
                with oring.begin_sequence(time_tag=iseq.time_tag, header=json.dumps(ohdr)) as oseq:
                    for ispan in iseq.read(igulp_size):
                        self.update_stats({'curr_sample': now})
                        if ispan.size < igulp_size:
                            continue
                        if self.update_pending:
                            with self._control_lock:
                                for b in range(self.nbeam):
                                    if self.gains_load_sample[b] == 0:
                                        continue
                                    if now >= self.gains_load_sample[b]:
                                        self.gains_cpu[:, b, :] = self.gains_cpu_new[:, b, :]
                                        self.gains_load_sample[b] = 0
                                        copy_pending = True
                                if self.gains_load_sample.sum() == 0:
                                    self.update_pending = False
                                self.stats['update_pending'] = self.update_pending
                                self.stats['last_cmd_proc_time'] = time.time()
                        if copy_pending:
                            self.log.debug("BEAMFORM >> Copy coefficients to GPU at time %d" % now)
                            rt.copy_array(self.gains_gpu, self.gains_cpu)
                            # the tensor-core kernel reads the weights as int8 digit planes: converted here, once per
                            # update, instead of with every gulp (bfBeamformRun takes them per call, :449)
                            self._bf.set_weights(self.gains_gpu)
                            copy_pending = False
                        t = time.time()
                        acquire_time, prev = t - prev, t
                        with oseq.reserve(ogulp_size) as ospan:
                            t = time.time()
                            reserve_time, prev = t - prev, t
                            if self.ntime_sum is None:
                                self._bf.run(ispan.data, ospan.data, None)
                            else:
                                self._bf.run_integrate(ispan.data, ospan.data, None, self.ntime_sum)
                            self._bf.sync()
                        now += self.ntime_gulp
                        t = time.time()
                        process_time, prev = t - prev, t
                        self.perf_proclog.update({'acquire_time': acquire_time, 'reserve_time': reserve_time,
                                                  'process_time': process_time,
                                                  'gbps': 8 * igulp_size / process_time / 1e9})
