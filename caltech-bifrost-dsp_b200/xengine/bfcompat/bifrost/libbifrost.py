"""``from bifrost.libbifrost import _bf, bf``: the reference's handle on the native library.  ``_bf`` is rebound to an
``XengBF`` (libxeng) by whoever assembles the pipeline -- ``bifrost.libbifrost.configure(nstand=..., nchan=...,
ntime_gulp=...)`` -- BEFORE the block modules are imported (they bind the name at import time)."""
from xengine.bfcompat.adapter import XengBF

_bf = XengBF()
bf = _bf


def configure(**sizes):
    """xGPU's compile-time constants (nstand, npol, nchan, ntime_gulp) for the process-wide ``_bf``."""
    _bf.cfg.update(sizes)
    return _bf
