"""Beamform launches at configs[3] size for `ncu --set full -k regex:beamform_tc_kernel` (tensor-core kernel, weights
pre-loaded: voltages out, then fused power integration)."""
import sys
sys.path.insert(0, "caltech-bifrost-dsp_b200")
import torch
from xengine import ops
nbeam, nt, nchan, ninput = 32, 960, 192, 704
bf = ops.Beamformer(ninput=ninput, nchan=nchan, ntime=nt, nbeam=nbeam)
vv = torch.randint(0, 256, (nt, nchan, ninput), dtype=torch.uint8, device="cuda")
w = torch.randn((nchan, nbeam, ninput, 2), dtype=torch.float32, device="cuda")
bo = torch.zeros((nchan, nbeam, nt, 2), dtype=torch.float32, device="cuda")
po = torch.zeros((nbeam // 2, nt // 24, nchan, 4), dtype=torch.float32, device="cuda")
bf.set_weights(w)
for _ in range(3):
    bf.run(vv, bo, None)
    bf.run_integrate(vv, po, None, 24)
bf.sync()
