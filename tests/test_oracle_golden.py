"""Pin the numpy oracle to fixtures produced by the reference's own code
(tests/golden/make_golden.py).  CPU only."""
import hashlib
import os

import numpy as np
import pytest

from oracle import xeng_oracle as orc


def _load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name), allow_pickle=False)


@pytest.mark.parametrize("tag", ["random", "chanramp"])
def test_corr_full_matches_reference_generator(golden_dir, tag):
    g = _load(golden_dir, "golden_corr_%s.npz" % tag)
    vin, acc = g["vin"], int(g["acc_len"])
    for blk in range(vin.shape[0] // acc):
        re, im = orc.corr_full(vin[blk * acc:(blk + 1) * acc])
        assert np.array_equal(re, g["corr_re"][blk])
        assert np.array_equal(im, g["corr_im"][blk])


def test_regtile_index_matches_reference(golden_dir):
    g = _load(golden_dir, "golden_regtile.npz")
    for nstand in (4, 8, 12):
        in0, in1, idx = orc.xgpu_maps(nstand)
        assert np.array_equal(idx, g["regtile_%d" % nstand])
    in0, in1, idx = orc.xgpu_maps(352)
    assert np.array_equal(idx[::997], g["regtile_352_sample"])
    assert idx.max() == int(g["regtile_352_max"])
    dig = np.frombuffer(hashlib.sha256(idx.astype("<i8").tobytes()).digest(), dtype=np.uint8)
    assert np.array_equal(dig, g["regtile_352_sha256"])
    tri = np.array([[orc.tri_index(i, j) for j in range(8)] for i in range(8)])
    assert np.array_equal(tri, g["tri"])


def test_regtile_352_geometry():
    """W = 249216 words/channel, 248512 used, 704 padding."""
    in0, in1, idx = orc.xgpu_maps(352)
    W = orc.words_per_chan(352)
    assert W == 249216 and W * 192 == 47849472     # xgpu_test.py:9
    assert idx.size == 248512 and np.unique(idx).size == idx.size
    assert idx.min() == 0 and idx.max() < W


def test_corr_xgpu_matches_Corr_test(golden_dir):
    """Corr._test (reference numpy correlator) read through regtile_index, as
    Corr._compare / xgpu_test.py do."""
    g = _load(golden_dir, "golden_corr_test.npz")
    vin = g["vin"]
    nt, nchan, nstand, npol = vin.shape
    x = orc.corr_xgpu(vin)
    lit = orc.corr_loop_literal(vin)
    assert np.array_equal(lit.real.astype(np.int32), g["out_re"])
    assert np.array_equal(lit.imag.astype(np.int32), g["out_im"])
    used = np.zeros(x.shape[2], dtype=bool)
    for i0 in range(nstand):
        for i1 in range(i0, nstand):
            for p in range(4):
                p0, p1 = p // 2, p % 2
                k = orc.regtile_index(2 * i0 + p0, 2 * i1 + p1, nstand)
                used[k] = True
                assert np.array_equal(x[0, :, k], g["out_re"][:, i0, i1, p])
                assert np.array_equal(x[1, :, k], g["out_im"][:, i0, i1, p])
    assert not x[:, :, ~used].any()          # padding words stay zero
    assert (~used).sum() == nstand // 2 * 4


def test_get_order_subselect_reorder_conventions(golden_dir):
    """bfXgpuGetOrder/SubSelect/Reorder net effect: v(s0,p0) conj(v(s1,p1)) in
    any stand order (corr_output_full_block.py:582-591, test_corr_part_rx.py:76-84)."""
    g = _load(golden_dir, "golden_corr_random.npz")
    vin, acc = g["vin"], int(g["acc_len"])
    v = vin[:acc]
    nt, nchan, nstand, npol = v.shape
    gold_re, gold_im = g["corr_re"][0], g["corr_im"][0]
    x = orc.corr_xgpu(v)
    a2i = np.arange(nstand * npol).reshape(nstand, npol)
    a2bl, conj = orc.get_order(a2i)
    ro = orc.reorder(x, a2bl, conj)
    assert np.array_equal(ro[..., 0], np.moveaxis(gold_re, 0, -1))
    assert np.array_equal(ro[..., 1], np.moveaxis(gold_im, 0, -1))
    # packet payloads of CorrOutputFull: one per baseline s0 <= s1, COR = [chan, pol, pol, 2], PY = [pol, pol, chan, 2]
    cor, py = orc.packet_payloads(ro, "cor"), orc.packet_payloads(ro, "py")
    assert cor.shape == (nstand * (nstand + 1) // 2, nchan, npol, npol, 2) and py.shape[1:] == (npol, npol, nchan, 2)
    bl = 0
    for s0 in range(nstand):
        for s1 in range(s0, nstand):
            assert np.array_equal(cor[bl, :, :, :, 0], gold_re[:, s0, s1])
            assert np.array_equal(cor[bl, :, :, :, 1], gold_im[:, s0, s1])
            assert py[bl].tobytes() == ro[s0, s1].tobytes()
            bl += 1
    rng = np.random.default_rng(1)
    bls = [[[int(a), int(b)], [int(c), int(d)]] for a, b, c, d in
           zip(rng.integers(0, nstand, 50), rng.integers(0, 2, 50),
               rng.integers(0, nstand, 50), rng.integers(0, 2, 50))]
    vm, cj = orc.baselines_to_vismap(bls, a2bl, conj)
    out = orc.subselect(x, vm, cj, nchan_sum=3)
    for n, ((s0, p0), (s1, p1)) in enumerate(bls):
        assert out[0, n, 0] == gold_re[:, s0, s1, p0, p1].sum()
        assert out[0, n, 1] == gold_im[:, s0, s1, p0, p1].sum()
    # CorrOutputPart payloads: plain = big-endian [vis, chan, 2]; COR = [vis/4, chan, 4, 2]
    sub = orc.subselect(x, vm[:48], cj[:48], nchan_sum=1)
    raw = orc.subsel_packet_payloads(sub, "py")
    assert np.array_equal(np.frombuffer(raw.tobytes(), dtype=">i4").reshape(48, nchan, 2), sub.transpose(1, 0, 2))
    corp = orc.subsel_packet_payloads(sub, "cor")
    for v in (0, 5, 47):
        assert np.array_equal(corp[v // 4, :, v % 4], sub[:, v])


def test_corr_accumulates_over_gulps(golden_dir):
    g = _load(golden_dir, "golden_corr_random.npz")
    vin = g["vin"]
    whole = orc.corr_xgpu(vin)
    acc = np.zeros(whole.shape, dtype=np.int64)
    for k in range(0, vin.shape[0], 16):
        orc.corr_xgpu(vin[k:k + 16], acc=acc)
    assert np.array_equal(acc, whole)
    a = np.zeros_like(whole)
    orc.corr_acc(a, whole, first=True)
    orc.corr_acc(a, whole, first=False)
    assert np.array_equal(a, 2 * whole)


def test_unpack_edge_bytes():
    re, im = orc.unpack4(np.array([0xFF, 0x88, 0x80, 0x08, 0x78, 0x87, 0x00], dtype=np.uint8))
    assert re.tolist() == [-1, -8, -8, 0, 7, -8, 0]
    assert im.tolist() == [-1, -8, 0, -8, -8, 7, 0]


def test_beamform_matches_reference(golden_dir):
    g = _load(golden_dir, "golden_beamform.npz")
    vin = g["vin"]
    nt, nchan, nstand, npol = vin.shape
    v = vin.reshape(nt, nchan, nstand * npol)
    r, i = orc.unpack4(v)
    assert np.array_equal((r + 1j * i).astype(np.complex64), g["unpacked"])
    out = orc.beamform(v, g["coeffs"])
    ref = g["out"]
    assert np.allclose(out, ref, rtol=1e-4, atol=1e-4 * np.abs(ref).max())  # ref is complex64 sums
    # weights built by the reference block from control JSON
    freqs = float(g["chan_bw_hz"]) * np.arange(nchan)
    for b in range(g["coeffs"].shape[1]):
        w = orc.beam_weights(freqs, g["delays"][b], g["amps"][b], g["calgains"][:, b, :])
        bg = g["block_gains"][:, b, :]
        assert np.allclose(w, bg, rtol=1e-5, atol=1e-5 * np.abs(bg).max())
        assert np.allclose(w, g["coeffs"][:, b, :], rtol=2e-4, atol=1e-3)


def test_beam_power_sum_matches_reference(golden_dir):
    g = _load(golden_dir, "golden_beamsum.npz")
    out = orc.beam_power_sum(g["x"], int(g["ntime_sum"]))
    assert out.shape == g["out"].shape
    assert np.allclose(out, g["out"], rtol=1e-5, atol=1e-4)
    one = orc.beam_power_sum_single(g["x"], int(g["ntime_sum"]), 1)
    assert np.allclose(one, g["out"][1], rtol=1e-5, atol=1e-4)


def test_snap2_packet_and_assembly_oracle():
    """SNAP2 header packing follows the reference transmitter byte for byte (test_tx_vectors.py:103-109) and
    assembling what was packetised gives the voltages back, zeros where packets were lost."""
    import struct
    rng = np.random.default_rng(3)
    nt, nchan, ninput, ncp, npp = 4, 8, 32, 4, 16
    v = rng.integers(0, 256, size=(nt, nchan, ninput), dtype=np.uint8)
    pkts = []
    for t in range(nt):
        for cb in range(nchan // ncp):
            for pb in range(ninput // npp):
                pkts.append(orc.snap2_packet(50 + t, 99, npp, ninput, ncp, nchan, cb, 100 + cb * ncp, pb * npp,
                                             v[t, cb * ncp:(cb + 1) * ncp, pb * npp:(pb + 1) * npp]))
    h = struct.unpack('>QLHHHHLLL', pkts[5][:32])
    assert h == (51, 99, npp, ninput, ncp, nchan, 0, 100, npp) and len(pkts[5]) == 32 + ncp * npp
    out, ngood = orc.snap2_assemble(pkts, 50, nt, nchan, ninput, first_chan=100)
    assert ngood == len(pkts) and np.array_equal(out, v)
    out, ngood = orc.snap2_assemble(pkts[1:], 50, nt, nchan, ninput, first_chan=100)
    assert ngood == len(pkts) - 1 and not out[0, :ncp, :npp].any() and np.array_equal(out[1:], v[1:])
    out, ngood = orc.snap2_assemble(pkts, 51, 2, nchan, ninput, first_chan=100)     # a later, shorter window
    assert ngood == 2 * (nchan // ncp) * (ninput // npp) and np.array_equal(out, v[1:3])


def test_upchannelize_oracle_against_literal_dft():
    """The f4 oracle (numpy fft) against the textbook sum X[k] = sum_n x[n] exp(-2 pi i n k / N) the reference's
    bifrost fft block computes (lwa352-upchan-bf.py:96: forward, unnormalised, no shift), both layouts; the N = 2
    case is exactly (x0 + x1, x0 - x1); the weighted sum is linear in the weights."""
    rng = np.random.default_rng(11)
    for nfine in (2, 8, 32):
        v = rng.integers(0, 256, size=(2, nfine, 3, 6), dtype=np.uint8)
        r, i = orc.unpack4(v)
        x = r.astype(np.float64) + 1j * i.astype(np.float64)              # [frame][n][chan][input]
        n = np.arange(nfine)
        lit = np.einsum("kn,fnci->fcik", np.exp(-2j * np.pi * np.outer(n, n) / nfine), x)
        got = orc.upchannelize(v)
        assert got.shape == (2, 3, 6, nfine) and np.allclose(got, lit, atol=1e-9)
        assert np.array_equal(orc.upchannelize(v, "freq_merged"), got.transpose(0, 1, 3, 2))
        if nfine == 2:
            assert np.array_equal(got[..., 0], x[:, 0] + x[:, 1]) and np.array_equal(got[..., 1], x[:, 0] - x[:, 1])
        w1 = rng.standard_normal((2, 3, 6, nfine)) + 1j * rng.standard_normal((2, 3, 6, nfine))
        w2 = rng.standard_normal((2, 3, 6, nfine)) + 1j * rng.standard_normal((2, 3, 6, nfine))
        assert np.allclose(orc.upchan_beamform(v, w1 + 2 * w2), orc.upchan_beamform(v, w1) + 2 * orc.upchan_beamform(v, w2))
    w = orc.offline_beam_weights([50e6, 50.1e6], [0.0, 10.0, 4.0])
    assert w.shape == (2, 3) and np.allclose(w[:, 1], 1.0) and np.isclose(w[0, 0], np.exp(2j * np.pi * 50e6 * 10e-9))


def test_snap2_oracle_against_the_reference_transmitter(golden_dir):
    """Fixture produced by running the reference's own SNAP2 emulator, pipeline/test_transmitters/test_tx_vectors.py
    (tests/golden/make_golden.py snap2_packets: its socket replaced by a recorder): every packet the reference sends
    is byte-identical to orc.snap2_packet with the header fields of test_tx_vectors.py:100-106, and assembling the
    recorded packets gives back the vector file the emulator was fed."""
    g = np.load(os.path.join(golden_dir, "golden_snap2_packets.npz"))
    data, pkts = g["data"], g["packets"]
    ntime, nchan, nstand, npol = data.shape
    ncb, spp = int(g["nchan_blocks"]), int(g["nstand_per_pkt"])
    cpp = nchan // ncb
    k = 0
    for t in range(ntime):                      # emission order: seq, channel block, stand block
        for cb in range(ncb):
            for pb in range(nstand // spp):
                payload = data[t, cb * cpp:(cb + 1) * cpp, pb * spp:(pb + 1) * spp, :].reshape(cpp, spp * npol)
                want = orc.snap2_packet(t, 0, npol * spp, nstand * npol, cpp, nchan, cb, cb * cpp, pb * spp * npol, payload)
                assert pkts[k].tobytes() == want, (t, cb, pb)
                k += 1
    assert k == len(pkts)
    out, ngood = orc.snap2_assemble([p.tobytes() for p in pkts], 0, ntime, nchan, nstand * npol)
    assert ngood == len(pkts) and np.array_equal(out, data.reshape(ntime, nchan, nstand * npol))


def test_upchan_correlate_oracle_is_hermitian_and_linear_in_frames():
    """orc.upchan_correlate (parity unpinned by reference vectors: bifrost's correlate block is an absent submodule):
    Hermitian, diagonal = summed power, and the sum over two halves of the frames equals the whole."""
    rng = np.random.default_rng(5)
    X = rng.standard_normal((12, 3, 10)) + 1j * rng.standard_normal((12, 3, 10))
    C = orc.upchan_correlate(X)
    assert np.allclose(C, np.conj(C.transpose(0, 2, 1)))
    assert np.allclose(C[:, np.arange(10), np.arange(10)].real, (np.abs(X) ** 2).sum(axis=0))
    assert np.allclose(C, orc.upchan_correlate(X[:5]) + orc.upchan_correlate(X[5:]))
    assert np.allclose(C[1, 2, 7], (X[:, 1, 2] * np.conj(X[:, 1, 7])).sum())
