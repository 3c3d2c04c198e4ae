"""BeamformSumSingleBeam: beam-power integration of one dual-pol beam.

Drop-in for ``lwa352_pipeline.blocks.beamform_sum_single_beam_block.BeamformSumSingleBeam`` (reference:
pipeline/lwa352_pipeline/blocks/beamform_sum_single_beam_block.py:20-127): same constructor signature; output
header ``nstand = 1``, ``nbeam = 1``, ``nbit = 32``, ``complex = True``, ``acc_len``, ``ntime_block``,
``beam_id``, ``npol = 2``; output layout f32 [ntime_gulp/ntime_sum][nchan][4].
``_bf.bfBeamformIntegrateSingleBeam`` (:114) becomes ``xeng_bf_integrate_single``.
"""
from ._beam_power import BeamPowerBlock


class BeamformSumSingleBeam(BeamPowerBlock):
    def __init__(self, log, iring, oring, nchan_max=256, ntime_gulp=2500, ntime_sum=24, guarantee=True, core=-1,
                 gpu=-1, beam_id=0, etcd_client=None):
        super(BeamformSumSingleBeam, self).__init__(log, iring, oring, guarantee, core, etcd_client=etcd_client)
        self.beam_id = beam_id
        self.nchan_max = nchan_max
        self._init_common(ntime_gulp, ntime_sum, gpu)
        self.ogulp_size = self.ntime_blocks * self.nchan_max * 4 * 4
        self.oring.resize(self.ntime_gulp * self.nchan_max * 4 * 4)

    def _plan(self, ihdr):
        nchan = ihdr['nchan']
        ohdr = dict(ihdr)
        ohdr.update({'nstand': 1, 'nbeam': 1, 'nbit': 32, 'complex': True, 'acc_len': self.ntime_sum,
                     'ntime_block': self.ntime_blocks, 'beam_id': self.beam_id, 'npol': 2})
        return {'nchan': nchan, 'nbeam_in': ihdr['nbeam'], 'ohdr': ohdr, 'beam_id': self.beam_id,
                'igulp_size': self.ntime_gulp * nchan * ihdr['nbeam'] * 2 * ihdr['nbit'] // 8,
                'ogulp_size': self.ntime_blocks * nchan * 4 * 4,
                'ring_size': self.ntime_gulp * self.nchan_max * 4 * 4}
