"""Developer tool: time the fused Corr+CorrAcc kernel with parts of it switched off (XENG_CORR_ABLATE bit mask:
1 no global write, 2 no epilogue staging stores, 4 no MMA issue, 8 no operand stores by the unpack warps,
16 no nibble arithmetic, 32 no TMA loads, 64 unpack warps read global memory directly (LDG, no raw ring, no TMA),
128 half of the B-column work (what one CTA of a cta_group::2 pair would do), 256 nothing (the ablation
instantiation's own baseline)).  The mask is read once per handle, at xeng_corr_create.
Results of ablated runs are invalid by construction; only the timings matter."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "caltech-bifrost-dsp_b200"))
from xengine import ops  # noqa: E402

nchan, ntime = int(os.environ.get("NCHAN", 192)), int(os.environ.get("NTIME", 480))
masks = [int(x) for x in sys.argv[1:]] or [0, 1, 3, 4, 8, 12, 7, 11, 15]
v = [torch.randint(0, 256, (ntime, nchan, 704), dtype=torch.uint8, device="cuda") for _ in range(4)]
acc = torch.zeros((2, nchan, ops.words_per_chan(352)), dtype=torch.int32, device="cuda")
for m in masks:
    os.environ["XENG_CORR_ABLATE"] = str(m)
    xc = ops.XCorr(nstand=352, nchan=nchan, ntime_gulp=ntime)
    xc.set_impl(int(os.environ.get("IMPL", 0)))
    for mode in (False, True):       # accumulate (reduce) / first (store)
        for i in range(5):
            xc.accumulate(v[i % 4], acc, mode)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        n = 40
        for i in range(n):
            xc.accumulate(v[i % 4], acc, mode)
        e1.record()
        torch.cuda.synchronize()
        print("ablate=%2d %s  %.1f us/gulp" % (m, "store " if mode else "reduce", 1e3 * e0.elapsed_time(e1) / n), flush=True)
os.environ["XENG_CORR_ABLATE"] = "0"
