"""BeamformSumBeams: integrate all dual-pol beam powers of a gulp.

Drop-in for ``lwa352_pipeline.blocks.beamform_sum_beams_block.BeamformSumBeams`` (reference:
pipeline/lwa352_pipeline/blocks/beamform_sum_beams_block.py:19-258): same constructor signature; output header
``nbeam //= 2``, ``nbit = 32``, ``complex = True``, ``acc_len = ntime_sum``, ``npol = 2``; output layout
f32 [nbeam/2][ntime_gulp/ntime_sum][nchan][4].  ``_bf.bfBeamformIntegrate`` (:245) becomes ``xeng_bf_integrate``.
"""
from ._beam_power import BeamPowerBlock


class BeamformSumBeams(BeamPowerBlock):
    def __init__(self, log, iring, oring, nchan=256, ntime_gulp=2500, ntime_sum=24, guarantee=True, core=-1, gpu=-1,
                 etcd_client=None):
        super(BeamformSumBeams, self).__init__(log, iring, oring, guarantee, core, etcd_client=etcd_client)
        self.nchan = nchan
        self._init_common(ntime_gulp, ntime_sum, gpu)

    def _plan(self, ihdr):
        nchan, nbeam = ihdr['nchan'], ihdr['nbeam']
        assert nchan == self.nchan
        ohdr = dict(ihdr)
        ohdr.update({'nbeam': nbeam // 2, 'nbit': 32, 'complex': True, 'acc_len': self.ntime_sum, 'npol': 2})
        ogulp_size = self.ntime_blocks * nchan * (nbeam // 2) * 4 * 4
        return {'nchan': nchan, 'nbeam_in': nbeam, 'ohdr': ohdr, 'beam_id': None,
                'igulp_size': self.ntime_gulp * nchan * nbeam * 8,
                'ogulp_size': ogulp_size,
                # spans are small: size the ring by input time so that a few gulps are buffered
                'ring_size': ogulp_size * self.ntime_sum * 4}
