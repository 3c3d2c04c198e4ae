"""GPU parity of the offline upchannelise path against the float64 oracle, through the C-ABI.
Tolerance: max |gpu - cpu| / max |cpu| <= 1e-5 (the beamformer's bar)."""
import numpy as np
import pytest

from oracle import xeng_oracle as orc

pytestmark = pytest.mark.gpu
TOL = 1e-5


@pytest.fixture(scope="module")
def torch_cuda():
    import torch
    if not torch.cuda.is_available():
        pytest.fail("gpu-marked test run without a CUDA device")
    return torch


def _edge_bytes(v):
    v.reshape(-1)[:4] = [0xFF, 0x88, 0x80, 0x08]   # -1-1j, -8-8j, -8, -8j: bytes randint(0, 255) never or rarely makes
    return v


def _cplx(t):
    a = t.cpu().numpy()
    return a[..., 0] + 1j * a[..., 1]


@pytest.mark.parametrize("nfine", [2, 4, 8, 16, 32])
@pytest.mark.parametrize("layout", ["fine_last", "freq_merged"])
@pytest.mark.parametrize("ninput", [704, 70])
def test_upchan_fft_matches_oracle(torch_cuda, nfine, layout, ninput):
    torch = torch_cuda
    from xengine import ops
    rng = np.random.default_rng(nfine * 1000 + ninput)
    nframe, nchan = 3, 5
    v = _edge_bytes(rng.integers(0, 256, size=(nframe, nfine, nchan, ninput), dtype=np.uint8))
    want = orc.upchannelize(v, layout)
    shape = (nframe, nchan, ninput, nfine, 2) if layout == "fine_last" else (nframe, nchan, nfine, ninput, 2)
    out = torch.full(shape, float("nan"), dtype=torch.float32, device="cuda")
    ops.upchan_fft(torch.from_numpy(v).cuda(), out, nframe, nfine, nchan, ninput,
                   ops.UPCHAN_FINE_LAST if layout == "fine_last" else ops.UPCHAN_FREQ_MERGED)
    torch.cuda.synchronize()
    got = _cplx(out)
    assert np.isfinite(got).all()
    assert np.abs(got - want).max() / np.abs(want).max() <= TOL


def test_upchan_fft_tone_lands_in_one_bin(torch_cuda):
    """A constant series (the DummySource ramp byte = stand % 8, dummy_source_block.py:194-195) has all its power in
    fine bin 0: X[0] = nfine * x, every other bin exactly 0."""
    torch = torch_cuda
    from xengine import ops
    nframe, nfine, nchan, ninput = 2, 32, 3, 64
    v = np.broadcast_to((np.arange(ninput) % 8).astype(np.uint8), (nframe, nfine, nchan, ninput)).copy()
    out = torch.empty((nframe, nchan, ninput, nfine, 2), dtype=torch.float32, device="cuda")
    ops.upchan_fft(torch.from_numpy(v).cuda(), out, nframe, nfine, nchan, ninput)
    torch.cuda.synchronize()
    got = _cplx(out)
    assert np.array_equal(got[..., 0].imag, np.broadcast_to(nfine * (np.arange(ninput) % 8), got.shape[:-1]))
    assert np.array_equal(got[..., 0].real, np.zeros(got.shape[:-1]))
    assert np.abs(got[..., 1:]).max() == 0.0


@pytest.mark.parametrize("nfine,nbeam,ninput,nframe", [(32, 1, 704, 4), (32, 4, 704, 7), (32, 6, 100, 5), (16, 3, 704, 11),
                                                       (8, 2, 704, 1), (4, 9, 40, 6), (2, 5, 96, 4)])
def test_upchan_beamform_matches_oracle(torch_cuda, nfine, nbeam, ninput, nframe):
    torch = torch_cuda
    from xengine import ops
    rng = np.random.default_rng(nfine + 10 * nbeam + ninput)
    nchan = 3
    v = _edge_bytes(rng.integers(0, 256, size=(nframe, nfine, nchan, ninput), dtype=np.uint8))
    # weights built the way BfOfflineWeightsBlock does: cal gains x geometric phases (beamform_offline_block.py:126-134,218)
    freqs = 50e6 + np.arange(nchan * nfine) * (23925.78125 / nfine)
    w = np.empty((nbeam, nchan, ninput, nfine), dtype=np.complex64)
    for b in range(nbeam):
        ph = orc.offline_beam_weights(freqs, rng.uniform(0, 500, ninput), rng.uniform(0, 1, ninput))
        cal = rng.standard_normal((nchan, ninput, nfine)) + 1j * rng.standard_normal((nchan, ninput, nfine))
        w[b] = (ph.reshape(nchan, nfine, ninput).transpose(0, 2, 1) * cal).astype(np.complex64)
    want = orc.upchan_beamform(v, w)
    out = torch.full((nframe, nbeam, nchan, nfine, 2), float("nan"), dtype=torch.float32, device="cuda")
    ops.upchan_beamform(torch.from_numpy(v).cuda(), torch.view_as_real(torch.from_numpy(w)).cuda(), out, nframe, nfine,
                        nchan, ninput, nbeam)
    torch.cuda.synchronize()
    got = _cplx(out)
    assert np.isfinite(got).all()
    assert np.abs(got - want).max() / np.abs(want).max() <= TOL


def test_upchan_beamform_equals_fft_then_weighted_sum_on_device(torch_cuda):
    """Full-size property (15 frames x 96 ch x 704 inputs x 32 bins, lwa352-upchan-bf.py:86-87): the fused kernel
    equals xeng_upchan_fft followed by a float64 weighted sum of its output."""
    torch = torch_cuda
    from xengine import ops
    nframe, nfine, nchan, ninput, nbeam = 15, 32, 96, 704, 2
    g = torch.Generator(device="cuda").manual_seed(7)
    v = torch.randint(0, 256, (nframe, nfine, nchan, ninput), dtype=torch.uint8, device="cuda", generator=g)
    w = torch.randn((nbeam, nchan, ninput, nfine, 2), dtype=torch.float32, device="cuda", generator=g)
    X = torch.empty((nframe, nchan, ninput, nfine, 2), dtype=torch.float32, device="cuda")
    out = torch.empty((nframe, nbeam, nchan, nfine, 2), dtype=torch.float32, device="cuda")
    ops.upchan_fft(v, X, nframe, nfine, nchan, ninput)
    ops.upchan_beamform(v, w, out, nframe, nfine, nchan, ninput, nbeam)
    torch.cuda.synchronize()
    Xc = torch.view_as_complex(X).to(torch.complex128)
    wc = torch.view_as_complex(w).to(torch.complex128)
    want = torch.einsum("bcik,fcik->fbck", wc, Xc)
    got = torch.view_as_complex(out).to(torch.complex128)
    assert ((got - want).abs().max() / want.abs().max()).item() <= TOL
    # the two input halves are summed into the output with atomics; two addends commute, so reruns are bit-equal
    for _ in range(3):
        again = torch.full_like(out, float("nan"))
        ops.upchan_beamform(v, w, again, nframe, nfine, nchan, ninput, nbeam)
        torch.cuda.synchronize()
        assert torch.equal(again, out)


def test_upchan_argument_errors(torch_cuda):
    torch = torch_cuda
    from xengine import ops
    v = torch.zeros(64, dtype=torch.uint8, device="cuda")
    o = torch.zeros(64 * 2, dtype=torch.float32, device="cuda")
    with pytest.raises(RuntimeError):
        ops.upchan_fft(v, o, 1, 3, 1, 4)       # nfine not a supported power of two
    with pytest.raises(RuntimeError):
        ops.upchan_fft(v, o, 1, 4, 1, 4, 7)    # unknown layout
    with pytest.raises(RuntimeError):
        ops.upchan_beamform(v, None, o, 1, 4, 1, 4, 1)


def _lower_tiles_mask(ninput, tile=64):
    """Elements xeng_upchan_correlate writes: the 64 x 64 tiles on or below the diagonal."""
    i = np.arange(ninput)
    return (i[:, None] // tile) >= (i[None, :] // tile)


@pytest.mark.parametrize("shape", [
    (240, 6, 704),      # the imaging script's gulp: 480 samples at NUPCHAN = 2 (lwa352-upchan-imag.py:85-86), 6 of its 192 freqs
    (15, 5, 70),        # 32 fine channels, ragged inputs
    (1, 2, 64),         # a single frame
    (33, 3, 130),       # frames not a multiple of the chunk, inputs two tiles and a bit
])
def test_upchan_correlate_matches_oracle_and_accumulates(torch_cuda, shape):
    torch = torch_cuda
    from xengine import ops
    nframe, nfreq, ninput = shape
    rng = np.random.default_rng(nframe * 7 + ninput)
    nfine = 2
    # genuine upchannelised voltages: the FFT kernel's frequency-merged output feeds the correlator, as in the script
    v = _edge_bytes(rng.integers(0, 256, size=(2, nframe, nfine, (nfreq + 1) // 2, ninput), dtype=np.uint8))
    mask = _lower_tiles_mask(ninput)
    out = torch.full((nfreq, ninput, ninput, 2), float("nan"), dtype=torch.float32, device="cuda")
    want = np.zeros((nfreq, ninput, ninput), dtype=np.complex128)
    for g in range(2):                                  # two gulps of one integration: first, then accumulate
        nchan = v.shape[3]
        X = torch.empty((nframe, nchan * nfine, ninput, 2), dtype=torch.float32, device="cuda")
        ops.upchan_fft(torch.from_numpy(v[g]).cuda(), X, nframe, nfine, nchan, ninput, ops.UPCHAN_FREQ_MERGED)
        Xs = X[:, :nfreq].contiguous()                  # FrequencySelectBlock: the first nfreq merged channels
        ops.upchan_correlate(Xs, out, nframe, nfreq, ninput, first=(g == 0))
        want += orc.upchan_correlate(orc.upchannelize(v[g], "freq_merged").reshape(nframe, nchan * nfine, ninput)[:, :nfreq])
    torch.cuda.synchronize()
    got = _cplx(out)
    assert np.isfinite(got[:, mask]).all()
    assert np.isnan(got[:, ~mask].real).all()           # the rest of the upper triangle is not touched
    err = np.abs(got[:, mask] - want[:, mask]).max(axis=1) / np.abs(want[:, mask]).max(axis=1)
    assert err.max() <= TOL
    # Hermitian where both halves exist (diagonal tiles), real non-negative autocorrelations
    d = np.arange(ninput)
    assert (got[:, d, d].real >= 0).all() and (np.abs(got[:, d, d].imag) <= TOL * got[:, d, d].real + 1e-30).all()


def test_upchan_correlate_refuses_bad_arguments(torch_cuda):
    torch = torch_cuda
    from xengine import ops
    from xengine._lib import XengError
    x = torch.zeros(64, dtype=torch.float32, device="cuda")
    with pytest.raises(XengError):
        ops.upchan_correlate(x, x, 1, 0, 4)
    with pytest.raises(XengError):
        ops.upchan_correlate(x, None, 1, 1, 4)
