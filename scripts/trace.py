"""Developer tool: per-role phase times of CTA 0 of the Corr kernel (XENG_CORR_TRACE, ablation build).
Usage: XENG_CORR_ABLATE=<mask|128> python scripts/trace.py   (bit 128 changes nothing but selects the traced build)"""
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
first = bool(int(os.environ.get("FIRST", "0")))
for i in range(3):
    xc.accumulate(v[i % 2], acc, first)
torch.cuda.synchronize()
path = os.path.join(ROOT, "gpurun_out", "trace.bin")
os.environ["XENG_CORR_TRACE"] = path
xc.accumulate(v[0], acc, first)
torch.cuda.synchronize()
del os.environ["XENG_CORR_TRACE"]
t = np.fromfile(path, dtype=np.int64)[:40].reshape(5, 8)
names = {
    0: ("unpack0", ["loop", "wait raw_full", "LDS+ALU", "wait empty", "STS", "fence", "-", "-"]),
    1: ("unpack1", ["loop", "wait raw_full", "LDS+ALU", "wait empty", "STS", "fence", "-", "-"]),
    2: ("mma", ["tile loop", "wait tmem_empty", "stage loop", "wait full", "issue+commit", "-", "-", "-"]),
    3: ("tma", ["issue", "wait raw_empty", "-", "-", "-", "-", "-", "-"]),
    4: ("epi w0", ["tile loop", "wait tmem_full", "-", "-", "ld+wait_read", "sts+ld", "fence", "bulk issue"]),
}
ntile = 28
for r in range(5):
    nm, ph = names[r]
    tot = int(t[r].sum())
    print("== %-8s total %8d cycles (%6.0f per tile)" % (nm, tot, tot / ntile))
    for i in range(8):
        if t[r, i]:
            print("     %-16s %8d  %5.1f%%  %7.0f /tile" % (ph[i], t[r, i], 100.0 * t[r, i] / tot, t[r, i] / ntile))
