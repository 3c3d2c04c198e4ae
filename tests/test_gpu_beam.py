"""GPU parity of Beamform / BeamformSum against the float64 numpy oracle, through the C-ABI.
Tolerance: north_star asks for 1e-5 relative in fp32 (the reference's own tests use 1e-4:
beamformer_test.py:109, beamformer_sum_test.py:104, both np.isclose(rtol=atol=1e-4)).
Relative error is normalised PER OUTPUT ROW -- per (channel, beam) for voltages, per (dual-pol beam, channel) for
powers: max_t |gpu - cpu| / max_t |cpu| -- which is stricter than one global max-norm (a weak beam is not hidden
behind a strong one).  A per-element ratio is not meaningful: a sum of 704 complex terms can cancel to ~0.

Every beamformer test runs on both implementations: impl 0 = tcgen05 int8 digit-plane kernel (csrc/xeng_beam_tc.cu;
shapes it cannot take fall through to the fp32 kernel inside the library), impl 1 = fp32 FMA kernel."""
import os

import numpy as np
import pytest

from oracle import xeng_oracle as orc

pytestmark = pytest.mark.gpu
RTOL = 1e-5


@pytest.fixture(scope="module")
def torch_cuda():
    import torch
    if not torch.cuda.is_available():
        pytest.fail("gpu-marked test run without a CUDA device")
    return torch


def _weights(rng, nchan, nbeam, ninput):
    """As beamformer_test.py:131-139 builds them: amps * exp(2 pi i f delay) * cal_gains."""
    freqs = 30e6 + 23925.78125 * np.arange(nchan)
    w = np.empty((nchan, nbeam, ninput), dtype=np.complex64)
    for b in range(nbeam):
        delays = rng.uniform(0, 1000, size=ninput)
        amps = rng.uniform(0, 1, size=ninput)
        cal = (rng.standard_normal((nchan, ninput)) + 1j * rng.standard_normal((nchan, ninput))).astype(np.complex64)
        w[:, b, :] = orc.beam_weights(freqs, delays, amps, cal)
    return w


def _beamform(torch, v, w, impl=0, preload=False):
    from xengine import ops
    nt, nchan, ninput = v.shape
    nbeam = w.shape[1]
    bf = ops.Beamformer(ninput=ninput, nchan=nchan, ntime=nt, nbeam=nbeam)
    bf.set_impl(impl)
    dv = torch.from_numpy(v).cuda()
    dw = torch.view_as_real(torch.from_numpy(w)).cuda()
    out = torch.full((nchan, nbeam, nt, 2), float("nan"), dtype=torch.float32, device="cuda")
    if preload:                 # weights converted once, run() without them
        bf.set_weights(dw)
        bf.run(dv, out, None)
    else:                       # the reference's calling convention: weights with every call
        bf.run(dv, out, dw)
    bf.sync()
    res = out.cpu().numpy()
    bf.close()
    return res[..., 0] + 1j * res[..., 1]


def row_err(got, want, axis=-1):
    """max over `axis` of |got - want| / max over `axis` of |want|, worst row."""
    num = np.abs(got - want).max(axis=axis)
    den = np.abs(want).max(axis=axis)
    den = np.where(den == 0, 1.0, den)
    return float((num / den).max())


IMPLS = [0, 1]


@pytest.mark.parametrize("shape", [
    (120, 16, 80, 32),      # beamformer_test.py: 40 stands x 2 pol, 16 ch, 120-sample gulps, 32 beams
    (480, 4, 704, 16),      # production inputs, config 4 beams
    (33, 3, 70, 5),         # ragged everything
    (1, 1, 1, 1),
    (130, 2, 704, 34),
])
@pytest.mark.parametrize("impl", IMPLS)
def test_beamform_matches_oracle(torch_cuda, shape, impl):
    nt, nchan, ninput, nbeam = shape
    rng = np.random.default_rng(0xaabbccdd)
    v = rng.integers(0, 256, size=(nt, nchan, ninput), dtype=np.uint8)
    v.reshape(-1)[:1] = 0x88
    w = _weights(rng, nchan, nbeam, ninput)
    got = _beamform(torch_cuda, v, w, impl)
    want = orc.beamform(v, w)
    assert np.isfinite(got).all()
    err = row_err(got, want)
    assert err <= RTOL, err


@pytest.mark.parametrize("shape", [(256, 3, 704, 32), (200, 2, 96, 8), (300, 2, 2032, 16), (64, 2, 64, 40)])
def test_beamform_tensor_core_against_fp32_kernel_and_preloaded_weights(torch_cuda, shape):
    """The two kernels are independent implementations (int8 digit planes on tcgen05 / fp32 FMA): they must agree to
    fp32 rounding, and set_weights + run(None) must equal run(weights) bit for bit."""
    nt, nchan, ninput, nbeam = shape
    rng = np.random.default_rng(nt + ninput)
    v = rng.integers(0, 256, size=(nt, nchan, ninput), dtype=np.uint8)
    w = _weights(rng, nchan, nbeam, ninput)
    tc = _beamform(torch_cuda, v, w, 0)
    fp = _beamform(torch_cuda, v, w, 1)
    pre = _beamform(torch_cuda, v, w, 0, preload=True)
    want = orc.beamform(v, w)
    assert row_err(tc, want) <= RTOL and row_err(fp, want) <= RTOL
    assert row_err(tc, fp) <= 5e-6
    assert np.array_equal(tc, pre)


def test_beamform_weight_dynamic_range(torch_cuda):
    """Block-floating-point digits are per (channel, beam) row: rows of wildly different scale (1e-20 .. 1e20), a row
    with one dominant weight, an all-zero row and exact powers of two must all come out to 1e-5 of their own row."""
    nt, nchan, ninput, nbeam = 128, 2, 704, 8
    rng = np.random.default_rng(5)
    v = rng.integers(0, 256, size=(nt, nchan, ninput), dtype=np.uint8)
    w = (rng.standard_normal((nchan, nbeam, ninput)) + 1j * rng.standard_normal((nchan, nbeam, ninput)))
    for b, sc in enumerate([1e-20, 1e20, 1.0, 3e-5, 7e9, 1.0, 1.0, 1.0]):
        w[:, b] *= sc
    w[:, 5, 17] = 1e4                  # one dominant weight: the others keep 24 - 13 bits relative to it
    w[:, 6] = 0
    w[:, 7] = 2.0 ** rng.integers(-3, 4, size=(nchan, ninput))
    w = w.astype(np.complex64)
    want = orc.beamform(v, w)
    for impl in IMPLS:
        got = _beamform(torch_cuda, v, w, impl)
        assert np.isfinite(got).all()
        assert row_err(got, want) <= RTOL, impl
        assert not got[:, 6].any()
    assert np.array_equal(_beamform(torch_cuda, v, w, 0)[:, 7], want[:, 7])     # power-of-two weights: exact


@pytest.mark.parametrize("impl", IMPLS)
def test_beamform_config3_full_size(torch_cuda, impl):
    """BASELINE configs[3] at its full size: 32 beams x 704 inputs x 192 channels x 960 samples, every channel
    against the float64 oracle."""
    nt, nchan, ninput, nbeam = 960, 192, 704, 32
    rng = np.random.default_rng(0xaabbccdd)
    v = rng.integers(0, 256, size=(nt, nchan, ninput), dtype=np.uint8)
    w = _weights(rng, nchan, nbeam, ninput)
    got = _beamform(torch_cuda, v, w, impl)
    want = orc.beamform(v, w)
    assert row_err(got, want) <= RTOL


@pytest.mark.parametrize("impl", IMPLS)
def test_beamform_golden(torch_cuda, golden_dir, impl):
    """Fixture generated by running the reference's own SoftwareBf arithmetic (tests/golden/make_golden.py)."""
    g = np.load(os.path.join(golden_dir, "golden_beamform.npz"))
    v, w, ref = g["vin"], np.ascontiguousarray(g["coeffs"].astype(np.complex64)), g["out"]
    v = np.ascontiguousarray(v.reshape(v.shape[0], v.shape[1], -1))
    got = _beamform(torch_cuda, v, w, impl)
    want = orc.beamform(v, w)
    assert row_err(got, want) <= RTOL
    assert row_err(got, ref) <= 1e-4   # the reference output itself is complex64 sums


@pytest.mark.parametrize("impl", IMPLS)
def test_beamform_kernel_test_pattern(torch_cuda, impl):
    """beamformer_kernel_test.py:29-31,45-46: data ((t%8)<<4), beam 0 weights 1, others 0."""
    nt, nchan, ninput, nbeam = 480, 6, 704, 16
    v = np.zeros((nt, nchan, ninput), dtype=np.uint8)
    v[...] = ((np.arange(nt) % 8) << 4)[:, None, None]
    w = np.zeros((nchan, nbeam, ninput), dtype=np.complex64)
    w[:, 0, :] = 1
    got = _beamform(torch_cuda, v, w, impl)
    assert np.array_equal(got, orc.beamform(v, w))   # small integers: exact
    assert not got[:, 1:].any()


def _integrate(torch, x, ntime_sum, single=None):
    from xengine import ops
    nchan, nbeam, nt = x.shape
    bf = ops.Beamformer(ninput=8, nchan=nchan, ntime=nt, nbeam=nbeam)
    dx = torch.view_as_real(torch.from_numpy(x.astype(np.complex64))).cuda()
    nblk = nt // ntime_sum
    if single is None:
        out = torch.zeros((nbeam // 2, nblk, nchan, 4), dtype=torch.float32, device="cuda")
        bf.integrate(dx, out, ntime_sum)
    else:
        out = torch.zeros((nblk, nchan, 4), dtype=torch.float32, device="cuda")
        bf.integrate_single(dx, out, ntime_sum, single)
    bf.sync()
    res = out.cpu().numpy()
    bf.close()
    return res


@pytest.mark.parametrize("shape", [(16, 32, 480, 24), (192, 32, 960, 24), (3, 2, 50, 7), (5, 4, 2200, 1100), (2, 6, 96, 1)])
def test_beam_power_sum(torch_cuda, shape):
    nchan, nbeam, nt, ns = shape
    rng = np.random.default_rng(0xaabbccdd)
    x = (rng.standard_normal((nchan, nbeam, nt)) + 1j * rng.standard_normal((nchan, nbeam, nt))).astype(np.complex64)
    got = _integrate(torch_cuda, x, ns)
    want = orc.beam_power_sum(x, ns)
    assert got.shape == want.shape
    assert np.abs(got - want).max() / np.abs(want).max() <= RTOL
    for b in range(nbeam // 2):
        one = _integrate(torch_cuda, x, ns, single=b)
        assert np.array_equal(one, got[b])


def test_beam_power_golden(torch_cuda, golden_dir):
    g = np.load(os.path.join(golden_dir, "golden_beamsum.npz"))
    x, ns, ref = np.ascontiguousarray(g["x"]), int(g["ntime_sum"]), g["out"]
    got = _integrate(torch_cuda, x, ns)
    want = orc.beam_power_sum(x, ns)
    assert np.abs(got - want).max() / np.abs(want).max() <= RTOL
    assert np.abs(got - ref).max() / np.abs(ref).max() <= 1e-4


def test_beamform_then_power_chain(torch_cuda):
    """Beamform -> BeamformSumBeams as production wires them (lwa352-pipeline.py:279-285)."""
    from xengine import ops
    torch = torch_cuda
    rng = np.random.default_rng(1)
    nt, nchan, ninput, nbeam, ns = 96, 8, 704, 8, 24
    v = rng.integers(0, 256, size=(nt, nchan, ninput), dtype=np.uint8)
    w = _weights(rng, nchan, nbeam, ninput)
    bf = ops.Beamformer(ninput=ninput, nchan=nchan, ntime=nt, nbeam=nbeam)
    volt = torch.zeros((nchan, nbeam, nt, 2), dtype=torch.float32, device="cuda")
    powr = torch.zeros((nbeam // 2, nt // ns, nchan, 4), dtype=torch.float32, device="cuda")
    bf.run(torch.from_numpy(v).cuda(), volt, torch.view_as_real(torch.from_numpy(w)).cuda())
    bf.integrate(volt, powr, ns)
    bf.sync()
    want = orc.beam_power_sum(orc.beamform(v, w), ns)
    got = powr.cpu().numpy()
    assert np.abs(got - want).max() / np.abs(want).max() <= RTOL


@pytest.mark.parametrize("impl", IMPLS)
@pytest.mark.parametrize("shape", [
    (960, 4, 704, 32, 24),    # configs[3] shape (fewer channels): 120-sample CTAs, 8 along time
    (960, 192, 704, 32, 24),  # configs[3] at full size
    (96, 8, 704, 8, 24),
    (252, 3, 70, 6, 7),       # ragged: 36 integration blocks, the last CTA partly empty, 6 beams in a 32-beam CTA
    (252, 3, 80, 6, 7),       # the same on the tensor-core path (ninput % 16 == 0)
    (128, 2, 64, 2, 128),     # one block per CTA
    (520, 2, 36, 4, 130),     # ntime_sum > 128: the two-kernel chain inside the library
    (384, 2, 128, 36, 12),    # two beam blocks on the tensor-core path
])
def test_beamform_fused_power(torch_cuda, shape, impl):
    """Beamform's ntime_sum mode (xeng_bf_run_integrate): power integration in the beamformer's epilogue equals
    the oracle within 1e-5 and equals the two-kernel chain (tap on) bit for bit."""
    from xengine import ops
    torch = torch_cuda
    nt, nchan, ninput, nbeam, ns = shape
    rng = np.random.default_rng(nt + ns)
    v = rng.integers(0, 256, size=(nt, nchan, ninput), dtype=np.uint8)
    w = _weights(rng, nchan, nbeam, ninput)
    dv, dw = torch.from_numpy(v).cuda(), torch.view_as_real(torch.from_numpy(w)).cuda()
    bf = ops.Beamformer(ninput=ninput, nchan=nchan, ntime=nt, nbeam=nbeam, ntime_blocks=nt // ns)
    bf.set_impl(impl)
    fused = torch.full((nbeam // 2, nt // ns, nchan, 4), float("nan"), dtype=torch.float32, device="cuda")
    bf.run_integrate(dv, fused, dw, ns)
    bf.sync()
    chain = torch.full_like(fused, float("nan"))
    bf.set_tap(True)
    bf.run_integrate(dv, chain, dw, ns)
    bf.sync()
    want = orc.beam_power_sum(orc.beamform(v, w)[:, :, :(nt // ns) * ns], ns)
    got = fused.cpu().numpy()
    assert np.isfinite(got).all()
    npair, nblk = got.shape[0], got.shape[1]       # per (dual-pol beam, channel): over time blocks and components
    assert row_err(got.transpose(0, 2, 1, 3).reshape(npair, nchan, nblk * 4),
                   want.transpose(0, 2, 1, 3).reshape(npair, nchan, nblk * 4)) <= RTOL
    assert np.array_equal(got, chain.cpu().numpy())
    bf.close()


@pytest.mark.parametrize("shape", [(192, 32, 960, 2), (5, 3, 70, 3), (33, 4, 33, 1)])
def test_vlbi_tap_transpose(torch_cuda, shape):
    """BeamformVlbiOutput's tap: idata[:, 0:nsend, :].transpose([2,0,1]) (beamform_vlbi_output_block.py:260-264),
    a pure permutation: bit-exact."""
    from xengine import ops
    torch = torch_cuda
    nchan, nbeam, nt, nsend = shape
    rng = np.random.default_rng(nchan + nt)
    x = (rng.standard_normal((nchan, nbeam, nt)) + 1j * rng.standard_normal((nchan, nbeam, nt))).astype(np.complex64)
    bf = ops.Beamformer(ninput=8, nchan=nchan, ntime=nt, nbeam=nbeam)
    out = torch.zeros((nt, nchan, nsend, 2), dtype=torch.float32, device="cuda")
    bf.vlbi_tap(torch.view_as_real(torch.from_numpy(x)).cuda(), out, nsend)
    bf.sync()
    got = out.cpu().numpy()
    want = np.ascontiguousarray(x[:, 0:nsend, :].transpose([2, 0, 1]))
    assert np.array_equal(got[..., 0] + 1j * got[..., 1], want)
    with pytest.raises(Exception):
        bf.vlbi_tap(out, out, nbeam + 1)
    bf.close()
