"""Two launches of each offline-upchannelise kernel at the scripts' sizes (15 frames x 32 fine x 96 ch x 704 inputs,
4 beams) for an ncu pass recording duration and DRAM bytes per launch (scripts/gpu_round.sh)."""
import sys
sys.path.insert(0, "caltech-bifrost-dsp_b200")
import torch
from xengine import ops

dev = "cuda"
nfr, nf, nch, nin, nb = 15, 32, 96, 704, 4
v = torch.randint(0, 256, (nfr, nf, nch, nin), dtype=torch.uint8, device=dev)
x = torch.zeros((nfr, nch, nin, nf, 2), dtype=torch.float32, device=dev)
w = torch.randn((nb, nch, nin, nf, 2), dtype=torch.float32, device=dev)
o = torch.zeros((nfr, nb, nch, nf, 2), dtype=torch.float32, device=dev)
for _ in range(2):
    ops.upchan_fft(v, x, nfr, nf, nch, nin, ops.UPCHAN_FINE_LAST)
    ops.upchan_fft(v, x, nfr, nf, nch, nin, ops.UPCHAN_FREQ_MERGED)
    ops.upchan_beamform(v, w, o, nfr, nf, nch, nin, nb)
    torch.cuda.synchronize()
print("ok")
