"""BeamformSum: beam-power integration for up to ``nbeam_max`` single-pol beams, written in place.

Drop-in for ``lwa352_pipeline.blocks.beamform_sum_block.BeamformSum`` (reference: pipeline/lwa352_pipeline/blocks/
beamform_sum_block.py:40-128): same constructor signature; output header ``nstand = 1``, ``nbeam = nbeam_max//2``,
``npol = 2``, ``nbit = 32``, ``complex = True``, ``acc_len``, ``ntime_block``.  The reference reserves
``ntime_blocks*nchan_max*nbeam_max*16`` bytes per gulp, twice what the powers occupy (:69); the span size is
kept for layout compatibility and the powers fill its first half.
"""
from ._beam_power import BeamPowerBlock


class BeamformSum(BeamPowerBlock):
    def __init__(self, log, iring, oring, tuning=0, nchan_max=256, nbeam_max=1, nstand=352, npol=2, ntime_gulp=2500,
                 ntime_sum=24, guarantee=True, core=-1, gpu=-1, etcd_client=None):
        super(BeamformSum, self).__init__(log, iring, oring, guarantee, core, etcd_client=etcd_client)
        self.tuning = tuning
        self.nchan_max, self.nbeam_max = nchan_max, nbeam_max
        self.nstand, self.npol, self.ninputs = nstand, npol, nstand * npol
        self._init_common(ntime_gulp, ntime_sum, gpu)

    def _plan(self, ihdr):
        ohdr = dict(ihdr)
        ohdr.update({'nstand': 1, 'nbeam': self.nbeam_max // 2, 'npol': 2, 'nbit': 32, 'complex': True,
                     'acc_len': self.ntime_sum, 'ntime_block': self.ntime_blocks})
        ogulp_size = self.ntime_blocks * self.nchan_max * self.nbeam_max * 4 * 4
        return {'nchan': self.nchan_max, 'nbeam_in': self.nbeam_max, 'ohdr': ohdr, 'beam_id': None,
                'igulp_size': self.ntime_gulp * self.nchan_max * self.nbeam_max * 8,
                'ogulp_size': ogulp_size, 'ring_size': ogulp_size}
