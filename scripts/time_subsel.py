"""Where does the fused fast dump's extra time go?  Fused Corr+CorrAcc(+CorrSubsel) at 704 x NCHAN x 480 with:
no selection kernel (plain), the SUB kernel with 4704 random / 4704 real (48 stands x all pols) / 48 words selected,
with and without the per-step zeroing of the fast buffer."""
import json
import sys
sys.path.insert(0, "caltech-bifrost-dsp_b200")
import numpy as np
import torch
from xengine import ops

nchan = int(sys.argv[1]) if len(sys.argv) > 1 else 192
nstand, ntime = 352, 480
dev = torch.device("cuda")
wpc = ops.words_per_chan(nstand)
gulps = [torch.randint(0, 256, (ntime, nchan, 2 * nstand), dtype=torch.uint8, device=dev) for _ in range(4)]
acc = torch.zeros(2 * nchan * wpc, dtype=torch.int32, device=dev)
a2bl, conj = ops.get_order(np.arange(2 * nstand, dtype=np.int32).reshape(nstand, 2))
xc = ops.XCorr(nstand=nstand, nchan=nchan, ntime_gulp=ntime)
k = [0]


def timeit(fn, n=30):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(n):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return float(np.median(ts))


def plain():
    xc.accumulate(gulps[k[0] % 4], acc, False)
    k[0] += 1


res = {"nchan": nchan, "plain": timeit(plain)}
rng = np.random.default_rng(1)
sels = {"random4704": rng.integers(0, a2bl.size, 4704)}
st = rng.choice(nstand, 48, replace=False)
real = [np.ravel_multi_index((s0, s1, p0, p1), a2bl.shape) for i, s0 in enumerate(st) for s1 in st[:i + 1]
        for p0 in range(2) for p1 in range(2)]
sels["stands48_allpol_%d" % len(real)] = np.array(real)
sels["autos48"] = np.array([np.ravel_multi_index((i, i, 0, 0), a2bl.shape) for i in range(48)])
for name, sel in sels.items():
    nv = len(sel)
    xc.set_subsel(a2bl.reshape(-1)[sel].copy(), conj.reshape(-1)[sel].copy(), 4)
    fast = torch.zeros((nchan // 4, nv, 2), dtype=torch.int32, device=dev)

    def fused_zero():
        xc.accumulate_subsel(gulps[k[0] % 4], acc, False, fast, True)
        k[0] += 1

    def fused_nozero():
        xc.accumulate_subsel(gulps[k[0] % 4], acc, False, fast, False)
        k[0] += 1
    res[name + "_zeroed_each_step"] = timeit(fused_zero)
    res[name + "_accumulating"] = timeit(fused_nozero)
res["plain_again"] = timeit(plain)
print(json.dumps(res, indent=1))
