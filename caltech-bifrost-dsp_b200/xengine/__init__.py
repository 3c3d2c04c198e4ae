"""B200-native LWA-352 X-engine hot path (Corr, CorrAcc, CorrSubsel, Beamform, BeamformSum) behind the
caltech-bifrost-dsp block API.  The arithmetic lives in libxeng.so (hand-written CUDA for sm_100a)."""
from . import _lib, ops  # noqa: F401

__all__ = ["_lib", "ops"]
