"""``bifrost.linalg.LinAlg``: imported by block_base.py but used by no X-engine block."""


class LinAlg(object):
    def matmul(self, *a, **k):
        raise NotImplementedError("bifrost.linalg is not part of the X-engine hot path")
