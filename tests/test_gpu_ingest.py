"""GPU parity of the SNAP2 packet -> voltage-buffer assembly against the oracle, through the C-ABI, and
end to end: packets -> xeng_snap2_assemble -> Corr equals the oracle correlation of the voltages that were sent."""
import numpy as np
import pytest

from oracle import xeng_oracle as orc

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def torch_cuda():
    import torch
    if not torch.cuda.is_available():
        pytest.fail("gpu-marked test run without a CUDA device")
    return torch


def _packets(v, seq0, first_chan, nchan_pkt, npol_pkt, sync_time=0xaabbccdd):
    """Packetise u8 [ntime][nchan][ninput] the way test_tx_vectors.py does: per spectrum, channel blocks outer,
    input ("pol") blocks inner."""
    nt, nchan, ninput = v.shape
    pkts = []
    for t in range(nt):
        for cb in range(nchan // nchan_pkt):
            for pb in range(ninput // npol_pkt):
                pkts.append(orc.snap2_packet(seq0 + t, sync_time, npol_pkt, ninput, nchan_pkt, nchan, cb,
                                             first_chan + cb * nchan_pkt, pb * npol_pkt,
                                             v[t, cb * nchan_pkt:(cb + 1) * nchan_pkt, pb * npol_pkt:(pb + 1) * npol_pkt]))
    return pkts


def _run(torch, pkts, stride, seq0, nt, nchan, ninput, first_chan, zero_first=True, out=None):
    from xengine import ops
    raw = np.zeros(len(pkts) * stride, dtype=np.uint8)
    for k, p in enumerate(pkts):
        raw[k * stride:k * stride + len(p)] = np.frombuffer(p, dtype=np.uint8)
    d_raw = torch.from_numpy(raw).cuda()
    if out is None:
        out = torch.full((nt, nchan, ninput), 0x77, dtype=torch.uint8, device="cuda")
    nvalid = torch.full((1,), -5, dtype=torch.int32, device="cuda")
    ops.snap2_assemble(d_raw, len(pkts), stride, out, seq0, nt, nchan, ninput, first_chan, zero_first, nvalid)
    torch.cuda.synchronize()
    return out, int(nvalid.cpu()[0])


@pytest.mark.parametrize("shape", [
    (8, 192, 704, 96, 64),     # production: 2 channel blocks x 11 input blocks per spectrum (test_tx.c:20-23)
    (5, 12, 40, 4, 8),         # 8-byte rows: the byte path
    (3, 6, 48, 6, 48),         # one packet per spectrum
])
def test_snap2_assemble_matches_oracle(torch_cuda, shape):
    nt, nchan, ninput, ncp, npp = shape
    rng = np.random.default_rng(sum(shape))
    v = rng.integers(0, 256, size=(nt, nchan, ninput), dtype=np.uint8)
    seq0, first_chan = (1 << 40) + 12345, 1672
    pkts = _packets(v, seq0, first_chan, ncp, npp)
    order = rng.permutation(len(pkts))                      # UDP does not keep order
    lost = set(int(i) for i in order[:len(pkts) // 7])      # and drops some
    sent = [pkts[i] for i in order if int(i) not in lost]
    # strays that must be ignored: before / after the window, another pipeline's channels, inputs past the end
    sent.append(orc.snap2_packet(seq0 - 1, 0, npp, ninput, ncp, nchan, 0, first_chan, 0, v[0, :ncp, :npp]))
    sent.append(orc.snap2_packet(seq0 + nt, 0, npp, ninput, ncp, nchan, 0, first_chan, 0, v[0, :ncp, :npp]))
    sent.append(orc.snap2_packet(seq0, 0, npp, ninput, ncp, nchan, 0, first_chan + nchan, 0, v[0, :ncp, :npp]))
    sent.append(orc.snap2_packet(seq0, 0, npp, ninput, ncp, nchan, 0, first_chan, ninput, v[0, :ncp, :npp]))
    stride = (32 + ncp * npp + 15) // 16 * 16
    got, ngood = _run(torch_cuda, sent, stride, seq0, nt, nchan, ninput, first_chan)
    want, wgood = orc.snap2_assemble(sent, seq0, nt, nchan, ninput, first_chan)
    assert ngood == wgood == len(pkts) - len(lost)
    assert np.array_equal(got.cpu().numpy(), want)
    # odd stride (unaligned payloads) takes the byte path and gives the same buffer
    got2, _ = _run(torch_cuda, sent, stride + 3, seq0, nt, nchan, ninput, first_chan)
    assert np.array_equal(got2.cpu().numpy(), want)
    # zero_first=False only overwrites what arrives
    keep = torch_cuda.full((nt, nchan, ninput), 0x77, dtype=torch_cuda.uint8, device="cuda")
    got3, _ = _run(torch_cuda, sent, stride, seq0, nt, nchan, ninput, first_chan, zero_first=False, out=keep)
    arrived = np.zeros((nt, nchan, ninput), dtype=bool)
    for i in range(len(pkts)):
        if i not in lost:
            t, rem = divmod(i, (nchan // ncp) * (ninput // npp))
            cb, pb = divmod(rem, ninput // npp)
            arrived[t, cb * ncp:(cb + 1) * ncp, pb * npp:(pb + 1) * npp] = True
    assert np.array_equal(got3.cpu().numpy(), np.where(arrived, v, 0x77))


def test_packets_to_visibilities(torch_cuda):
    """Packets in, visibilities out: the assembled buffer feeds Corr directly (same layout as the gpu-input ring)."""
    from xengine import ops
    torch = torch_cuda
    nt, nchan, nstand = 32, 8, 16
    rng = np.random.default_rng(7)
    v = rng.integers(0, 256, size=(nt, nchan, nstand * 2), dtype=np.uint8)
    pkts = _packets(v, 1000, 0, 4, 16)
    buf, ngood = _run(torch, pkts, 32 + 4 * 16, 1000, nt, nchan, nstand * 2, 0)
    assert ngood == len(pkts)
    xc = ops.XCorr(nstand=nstand, nchan=nchan, ntime_gulp=nt)
    acc = torch.zeros((2, nchan, ops.words_per_chan(nstand)), dtype=torch.int32, device="cuda")
    xc.accumulate(buf, acc, True)
    xc.sync()
    assert np.array_equal(acc.cpu().numpy(), orc.corr_xgpu(v.reshape(nt, nchan, nstand, 2)))


def test_snap2_assemble_reference_transmitter_packets(torch_cuda, golden_dir):
    """The packets the reference's own emulator sends (golden_snap2_packets.npz, recorded from
    pipeline/test_transmitters/test_tx_vectors.py), shuffled, through xeng_snap2_assemble: the voltage buffer equals
    the emulator's input vectors."""
    import os
    from xengine import ops
    torch = torch_cuda
    g = np.load(os.path.join(golden_dir, "golden_snap2_packets.npz"))
    data, pkts = g["data"], g["packets"]
    ntime, nchan, nstand, npol = data.shape
    ninput = nstand * npol
    order = np.random.default_rng(3).permutation(len(pkts))
    raw = torch.from_numpy(np.ascontiguousarray(pkts[order])).cuda()
    out = torch.full((ntime, nchan, ninput), 0xAA, dtype=torch.uint8, device="cuda")
    nvalid = torch.zeros(1, dtype=torch.int32, device="cuda")
    ops.snap2_assemble(raw, len(pkts), pkts.shape[1], out, 0, ntime, nchan, ninput, 0, True, nvalid)
    torch.cuda.synchronize()
    assert int(nvalid) == len(pkts)
    assert np.array_equal(out.cpu().numpy(), data.reshape(ntime, nchan, ninput))
