"""One process holding X-engine handles on TWO GPUs (kernel attributes and __constant__ tables live in a
device's context, so a process-wide "configured" flag left the second device unconfigured).  Needs >= 2 devices:
run with `gpurun --gpus 2`; on a single-GPU box the test is skipped."""
import numpy as np
import pytest

from oracle import xeng_oracle as orc

pytestmark = pytest.mark.gpu


def test_handles_on_two_devices_in_one_process():
    import torch
    from xengine import ops
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two CUDA devices")
    rng = np.random.default_rng(2)
    nt, nchan, nstand = 96, 4, 64
    v = rng.integers(0, 256, size=(nt, nchan, nstand, 2), dtype=np.uint8)
    want = orc.corr_xgpu(v)
    ninput, nbeam = nstand * 2, 8
    w = (rng.standard_normal((nchan, nbeam, ninput)) + 1j * rng.standard_normal((nchan, nbeam, ninput))).astype(np.complex64)
    wantv = orc.beamform(v.reshape(nt, nchan, ninput), w)
    for dev in (0, 1, 0):                       # device 1 is configured second; device 0 again afterwards
        xc = ops.XCorr(nstand=nstand, nchan=nchan, ntime_gulp=nt, gpu=dev)
        out = torch.zeros((2, nchan, ops.words_per_chan(nstand)), dtype=torch.int32, device="cuda:%d" % dev)
        for impl in (0, 2, 1):                  # TMA-fed tcgen05, register-fed tcgen05, CUDA-core cross-check
            xc.set_impl(impl)
            out.fill_(-1)
            xc.run(torch.from_numpy(v).to("cuda:%d" % dev), out, True)
            xc.sync()
            assert np.array_equal(out.cpu().numpy(), want), (dev, impl)
        xc.close()
        bf = ops.Beamformer(gpu=dev, ninput=ninput, nchan=nchan, ntime=nt, nbeam=nbeam)
        for impl in (0, 1):
            bf.set_impl(impl)
            volt = torch.zeros((nchan, nbeam, nt, 2), dtype=torch.float32, device="cuda:%d" % dev)
            bf.run(torch.from_numpy(v.reshape(nt, nchan, ninput)).to("cuda:%d" % dev), volt,
                   torch.view_as_real(torch.from_numpy(w)).to("cuda:%d" % dev))
            bf.sync()
            g = volt.cpu().numpy()
            assert np.abs(g[..., 0] + 1j * g[..., 1] - wantv).max() / np.abs(wantv).max() <= 1e-5, (dev, impl)
        bf.close()


@pytest.mark.gpu
def test_peer_push_payload_then_sequence_word():
    """xeng_peer_push: the payload lands, then the sequence word; misaligned calls are refused."""
    import torch
    from xengine import ops
    from xengine._lib import XengError
    dev = torch.device("cuda:0")
    n = 4704 * 48 * 2 + 4                                     # a fast dump's words, plus a ragged tail of one int4
    src = torch.randint(-2 ** 31, 2 ** 31 - 1, (n,), dtype=torch.int32, device=dev)
    dst = torch.zeros(n + 8, dtype=torch.int32, device=dev)
    flag = torch.full((2,), -1, dtype=torch.int64, device=dev)
    stream = torch.cuda.Stream(device=dev)
    with torch.cuda.stream(stream):
        ops.peer_push(dst[4:], src, n * 4, flag[1:], 77, stream)
    stream.synchronize()
    assert torch.equal(dst[4:4 + n], src) and int(dst[:4].abs().sum()) == 0 and int(dst[4 + n:].abs().sum()) == 0
    assert flag.tolist() == [-1, 77]
    keep = src.clone()
    dst.zero_()
    ops.peer_move(dst[4:], src, n * 4, flag[1:], 78, stream)  # the moving form: source zeroed behind the copy
    stream.synchronize()
    assert torch.equal(dst[4:4 + n], keep) and int(src.abs().sum()) == 0 and flag.tolist() == [-1, 78]
    src.copy_(keep)
    flag[1] = 77
    ops.peer_push(dst, src, 0, flag, 5, stream)             # empty payload: only the word moves
    stream.synchronize()
    assert flag.tolist() == [5, 77]
    if torch.cuda.device_count() > 1 and torch.cuda.can_device_access_peer(0, 1):
        far = torch.zeros(n, dtype=torch.int32, device="cuda:1")
        far.copy_(src)                                        # torch enables peer access on its first peer copy
        far.zero_()
        torch.cuda.synchronize("cuda:1")
        with torch.cuda.device(dev):
            ops.peer_push(far, src, n * 4, None, 0, stream)
            stream.synchronize()
        assert torch.equal(far.to(dev), src)
    for bad in (lambda: ops.peer_push(dst[1:], src, 16, None, 0, stream),
                lambda: ops.peer_push(dst, src, 20, None, 0, stream)):
        with pytest.raises(XengError):
            bad()


def test_peer_exchanges_two_ranks():
    """PeerStream / PeerGather (torch symmetric memory over NVLink) between two ranks, one per GPU."""
    import os
    import subprocess
    import sys
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two CUDA devices")
    here = os.path.dirname(os.path.abspath(__file__))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29533", os.path.join(here, "peer_worker.py")]
    res = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, timeout=240)
    assert res.returncode == 0, res.stdout.decode("utf-8", "replace")[-4000:]
