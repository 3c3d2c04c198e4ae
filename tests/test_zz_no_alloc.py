"""xeng_beam_power_static: the no-allocation Beamform + BeamformSum entry point (include/xeng.h; no counterpart in the
reference), called through the C-ABI twice in a row and compared with the existing entry point
xeng_bf_run_integrate (fp32 kernel, and the default tensor-core kernel) and with the float64 oracle, at the beam tests'
tolerance (tests/test_gpu_beam.py: 1e-5 relative per output row).  Named test_zz_* so that pytest collects it last."""
import ctypes as C

import numpy as np
import pytest

from oracle import xeng_oracle as orc

pytestmark = pytest.mark.gpu
RTOL = 1e-5


def _row_err(got, want):
    """[nbeam/2][nblock][nchan][4] powers: worst (dual-pol beam, channel, product) row, normalised by the row's max."""
    num = np.abs(got - want).max(axis=1)
    den = np.abs(want).max(axis=1)
    den = np.where(den == 0, 1.0, den)
    return float((num / den).max())


def _case(rng, nt, nchan, ninput, nbeam):
    v = rng.integers(0, 256, size=(nt, nchan, ninput), dtype=np.uint8)
    w = (rng.standard_normal((nchan, nbeam, ninput)) + 1j * rng.standard_normal((nchan, nbeam, ninput))).astype(np.complex64)
    return v, w


def test_static_scratch_entry_point_matches_run_integrate_twice_in_a_row():
    import torch
    from xengine import _lib, ops
    if not torch.cuda.is_available():
        pytest.fail("gpu-marked test run without a CUDA device")
    lib = _lib.load()
    rng = np.random.default_rng(77)
    nsum = 24
    # second call: a smaller problem, so the array still holds the first call's voltages past its end
    for nt, nchan, ninput, nbeam in ((480, 8, 704, 16), (240, 3, 80, 6)):
        v, w = _case(rng, nt, nchan, ninput, nbeam)
        dv = torch.from_numpy(v).cuda()
        dw = torch.view_as_real(torch.from_numpy(w)).contiguous().cuda()
        shape = (nbeam // 2, nt // nsum, nchan, 4)
        got = torch.full(shape, float("nan"), dtype=torch.float32, device="cuda")
        stream = torch.cuda.current_stream()
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated()
        rc = lib.xeng_beam_power_static(_lib.ptr(dv), _lib.ptr(dw), _lib.ptr(got), ninput, nchan, nt, nbeam, nsum,
                                        _lib.ptr(stream))
        _lib.check(rc, "xeng_beam_power_static")
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated() == before
        got = got.cpu().numpy()
        assert np.isfinite(got).all()

        for impl in (1, 0):     # the existing entry point: fp32 kernel (same arithmetic), tensor-core kernel
            bf = ops.Beamformer(ninput=ninput, nchan=nchan, ntime=nt, nbeam=nbeam)
            bf.set_impl(impl)
            ref = torch.full(shape, float("nan"), dtype=torch.float32, device="cuda")
            bf.run_integrate(dv, ref, dw, nsum)
            bf.sync()
            assert _row_err(got, ref.cpu().numpy()) <= RTOL, impl
            bf.close()

        want = orc.beam_power_sum(orc.beamform(v, w), nsum)
        assert _row_err(got, want) <= RTOL


def test_static_scratch_entry_point_refuses_what_does_not_fit():
    import torch
    from xengine import _lib
    lib = _lib.load()
    one = torch.zeros(16, dtype=torch.uint8, device="cuda")
    # 192 ch x 32 beams x 960 samples = 5.9 M voltages > 2 Mi: refused before anything is launched or dereferenced
    rc = lib.xeng_beam_power_static(_lib.ptr(one), _lib.ptr(one), _lib.ptr(one), 704, 192, 960, 32, 24, C.c_void_p(0))
    assert rc == -1
    with pytest.raises(_lib.XengError, match="static scratch"):
        _lib.check(rc, "xeng_beam_power_static")
    rc = lib.xeng_beam_power_static(_lib.ptr(one), _lib.ptr(one), _lib.ptr(one), 704, 4, 100, 8, 24, C.c_void_p(0))
    assert rc == -1                       # ntime_sum does not divide ntime
