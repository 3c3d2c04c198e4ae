"""CUDA-event timing (L2 flushed) of the fused upchannelise+beamform kernel at the offline script's size."""
import sys
sys.path.insert(0, "caltech-bifrost-dsp_b200")
import numpy as np
import torch
from xengine import ops

dev = "cuda"
nfr, nf, nch, nin, nb = 15, 32, 96, 704, 4
v = torch.randint(0, 256, (nfr, nf, nch, nin), dtype=torch.uint8, device=dev)
w = torch.randn((nb, nch, nin, nf, 2), dtype=torch.float32, device=dev)
o = torch.zeros((nfr, nb, nch, nf, 2), dtype=torch.float32, device=dev)
flush = torch.zeros(256 << 20, dtype=torch.uint8, device=dev)
for nbb in (1, 4):
    ts = []
    for i in range(14):
        flush.add_(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ops.upchan_beamform(v, w, o, nfr, nf, nch, nin, nbb)
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    print("upchan_beamform nbeam", nbb, "ms", float(np.median(ts[2:])))
