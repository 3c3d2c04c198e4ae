"""``bifrost.unpack.unpack``: imported by beamform_block.py:10 and beamform_sum_beams_block.py, called by neither."""


def unpack(*a, **k):
    raise NotImplementedError("bifrost.unpack is not part of the X-engine hot path")
