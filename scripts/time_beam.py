"""CUDA-event timings of the beamformer at BASELINE configs[3] size (32 beams x 704 inputs x 192 ch x 960 samples):
tensor-core kernel with per-call weight conversion (the reference's calling convention), with pre-loaded weights,
fused with the power integration, the weight conversion alone, and the fp32 FMA kernel.  L2 flushed between launches."""
import json
import sys
sys.path.insert(0, "caltech-bifrost-dsp_b200")
import numpy as np
import torch
from xengine import ops

nbeam, nt, nchan, ninput, nsum = 32, 960, 192, 704, 24
dev = torch.device("cuda")
flush = torch.zeros(256 << 20, dtype=torch.uint8, device=dev)
vv = torch.randint(0, 256, (nt, nchan, ninput), dtype=torch.uint8, device=dev)
w = torch.randn((nchan, nbeam, ninput, 2), dtype=torch.float32, device=dev)
bo = torch.zeros((nchan, nbeam, nt, 2), dtype=torch.float32, device=dev)
po = torch.zeros((nbeam // 2, nt // nsum, nchan, 4), dtype=torch.float32, device=dev)


def timeit(fn, n=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(n):
        flush.add_(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return float(np.median(ts)), float(min(ts))


res = {}
for impl, name in ((0, "tc"), (1, "fp32")):
    bf = ops.Beamformer(ninput=ninput, nchan=nchan, ntime=nt, nbeam=nbeam)
    bf.set_impl(impl)
    res[name + "_run_with_weights"] = timeit(lambda: bf.run(vv, bo, w))
    bf.set_weights(w)
    res[name + "_run_preloaded"] = timeit(lambda: bf.run(vv, bo, None))
    res[name + "_run_integrate_preloaded"] = timeit(lambda: bf.run_integrate(vv, po, None, nsum))
    if impl == 0:
        res["weight_conversion"] = timeit(lambda: bf.set_weights(w))
    bf.close()
fl = 8.0 * nbeam * ninput * nchan * nt
out = {k: {"ms_median": v[0], "ms_min": v[1], "gsamples_per_s": nt * nchan * ninput / v[0] / 1e6,
           "fp32_equiv_tflops": fl / v[0] / 1e9} for k, v in res.items()}
print(json.dumps(out, indent=1))
