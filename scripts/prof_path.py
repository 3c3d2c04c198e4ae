"""Two launches of every bandwidth-bound path kernel at BASELINE sizes, for an ncu pass that records duration and DRAM
bytes per launch (scripts/gpu_round.sh): CorrAcc, CorrSubsel (ring and packet layouts), reorder (full / COR payloads),
Beamform, BeamformSum, fused Beamform+power."""
import sys
sys.path.insert(0, "caltech-bifrost-dsp_b200")
import numpy as np
import torch
from xengine import ops

NSTAND, NCHAN = 352, 192
dev = "cuda"
wpc = ops.words_per_chan(NSTAND)
nw = 2 * NCHAN * wpc
vis = torch.randint(-1000, 1000, (nw,), dtype=torch.int32, device=dev)
acc = torch.zeros(nw, dtype=torch.int32, device=dev)
a2bl, conj = ops.get_order(np.arange(NSTAND * 2, dtype=np.int32).reshape(NSTAND, 2))
da, dc = torch.from_numpy(a2bl).to(dev), torch.from_numpy(conj).to(dev)
rng = np.random.default_rng(1)
# CorrSubsel's selection shape (corr_subsel_block.py:185): 48 stands, all pairs, 4 pol products = whole 16-byte cells
st = rng.choice(NSTAND, 48, replace=False)
idx = [(s0, s1, p0, p1) for i, s0 in enumerate(st) for s1 in st[:i + 1] for p0 in range(2) for p1 in range(2)]
flat = np.array([np.ravel_multi_index(t, a2bl.shape) for t in idx])
vm = torch.from_numpy(a2bl.reshape(-1)[flat].copy()).to(dev)
cj = torch.from_numpy(conj.reshape(-1)[flat].copy()).to(dev)
sel = rng.integers(0, a2bl.size, 4704)          # and 4704 random single words: every word in a cell of its own
vm_r = torch.from_numpy(a2bl.reshape(-1)[sel].copy()).to(dev)
cj_r = torch.from_numpy(conj.reshape(-1)[sel].copy()).to(dev)
so = torch.zeros((NCHAN // 4) * 4704 * 2, dtype=torch.int32, device=dev)
ro = torch.zeros(NSTAND * NSTAND * 4 * NCHAN * 2, dtype=torch.int32, device=dev)
pk = torch.zeros(NSTAND * (NSTAND + 1) // 2 * NCHAN * 8, dtype=torch.int32, device=dev)
nbeam, nt, nsum = 32, 960, 24
bf = ops.Beamformer(ninput=704, nchan=NCHAN, ntime=nt, nbeam=nbeam)
vv = torch.randint(0, 256, (nt, NCHAN, 704), dtype=torch.uint8, device=dev)
w = torch.randn((NCHAN, nbeam, 704, 2), dtype=torch.float32, device=dev)
bo = torch.zeros((NCHAN, nbeam, nt, 2), dtype=torch.float32, device=dev)
po = torch.zeros((nbeam // 2, nt // nsum, NCHAN, 4), dtype=torch.float32, device=dev)
for _ in range(2):
    ops.acc(acc, vis, nw, False)
    ops.subsel(vis, so, vm, cj, 4704, NCHAN, 4, wpc)
    ops.subsel(vis, so, vm_r, cj_r, 4704, NCHAN, 4, wpc)
    ops.subsel_packets(vis, so, vm, cj, 4704, NCHAN, 4, wpc, ops.PAYLOAD_COR)
    ops.reorder(vis, ro, da, dc, NSTAND, 2, NCHAN, wpc)
    ops.reorder_packets(vis, pk, da, dc, NSTAND, 2, NCHAN, wpc, ops.PAYLOAD_COR)
    bf.run(vv, bo, w)                   # weight conversion + tensor-core beamformer
    bf.integrate(bo, po, nsum)
    bf.run_integrate(vv, po, None, nsum)  # fused power epilogue, weights already loaded
    bf.set_impl(1)
    bf.run(vv, bo, w)                   # fp32 FMA kernel (cross-check implementation)
    bf.set_impl(0)
    torch.cuda.synchronize()
