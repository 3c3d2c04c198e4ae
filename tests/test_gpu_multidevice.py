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
