"""Developer tool: dump the per-role clock64 timeline of CTA 0 of the Corr kernel (XENG_CORR_TRACE, ablation build).
Usage: XENG_CORR_ABLATE=<mask|128> python scripts/trace.py   (mask bit 128 changes nothing but selects the traced build)"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "caltech-bifrost-dsp_b200"))
from xengine import ops  # noqa: E402

nchan, ntime = 192, 480
v = [torch.randint(0, 256, (ntime, nchan, 704), dtype=torch.uint8, device="cuda") for _ in range(2)]
acc = torch.zeros((2, nchan, ops.words_per_chan(352)), dtype=torch.int32, device="cuda")
xc = ops.XCorr(nstand=352, nchan=nchan, ntime_gulp=ntime)
os.environ.setdefault("XENG_CORR_ABLATE", "128")
for i in range(3):
    xc.accumulate(v[i % 2], acc, False)
torch.cuda.synchronize()
path = os.path.join(ROOT, "gpurun_out", "trace.bin")
os.environ["XENG_CORR_TRACE"] = path
xc.accumulate(v[0], acc, False)
torch.cuda.synchronize()
del os.environ["XENG_CORR_TRACE"]
t = np.fromfile(path, dtype=np.int64).reshape(5, 4096, 2)
names = {0: "unpack0", 1: "unpack1", 2: "mma", 3: "tma", 4: "epi"}
t0 = min(int(t[r, 0, 1]) for r in range(5) if t[r, 0, 0])
for r in range(5):
    ev = t[r]
    n = int((ev[:, 0] != 0).sum())
    print("== %s: %d events" % (names[r], n))
    # mean duration between consecutive tags, over events 200.. (steady state)
    import collections
    d = collections.defaultdict(list)
    for i in range(min(200, n // 4), n - 1):
        d[(int(ev[i, 0]), int(ev[i + 1, 0]))].append(int(ev[i + 1, 1] - ev[i, 1]))
    for k in sorted(d):
        a = np.array(d[k])
        print("   %d->%d  n=%4d mean=%7.1f med=%7.1f max=%d" % (k[0], k[1], len(a), a.mean(), np.median(a), a.max()))
    if n:
        print("   span: %d cycles for %d events" % (int(ev[n - 1, 1] - ev[0, 1]), n))
