"""Per-kernel timings on one B200 (CUDA events on the launching stream, L2 flushed between iterations).
Developer tool; bench.py is the contract benchmark.  Usage: python scripts/microbench.py [nchan] [iters]"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "caltech-bifrost-dsp_b200"))
from xengine import ops  # noqa: E402


def timeit(fn, iters, flush):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        if flush is not None:
            flush.add_(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return float(np.median(ts)), float(np.min(ts))


def main():
    nchan = int(sys.argv[1]) if len(sys.argv) > 1 else 192
    iters = int(sys.argv[2]) if len(sys.argv) > 2 else 10
    ntime = int(sys.argv[3]) if len(sys.argv) > 3 else 480
    nstand, npol = 352, 2
    ninput = nstand * npol
    dev = "cuda"
    flush = torch.zeros(256 << 20, dtype=torch.uint8, device=dev)  # 256 MiB > 126 MB L2
    W = ops.words_per_chan(nstand)
    res = {"nchan": nchan, "ntime": ntime}
    v = torch.randint(0, 256, (ntime, nchan, ninput), dtype=torch.uint8, device=dev)
    out = torch.zeros((2, nchan, W), dtype=torch.int32, device=dev)
    acc = torch.zeros_like(out)
    xc = ops.XCorr(nstand=nstand, nchan=nchan, ntime_gulp=ntime)
    nsamp = ntime * nchan * ninput
    useful_ops = 8 * 248512 * nchan * ntime

    med, best = timeit(lambda: xc.run(v, out, True), iters, flush)
    res["corr_dump_ms"] = med
    res["corr_dump_gsps"] = nsamp / med / 1e6
    res["corr_dump_useful_tops"] = useful_ops / med / 1e9
    res["corr_dump_out_gbs"] = out.numel() * 4 / med / 1e6
    med, best = timeit(lambda: xc.accumulate(v, acc, False), iters, flush)
    res["corr_accum_ms"] = med
    res["corr_accum_gsps"] = nsamp / med / 1e6
    res["corr_accum_useful_tops"] = useful_ops / med / 1e9
    xc.set_impl(1)
    if nchan <= 16:
        med, best = timeit(lambda: xc.run(v, out, True), 2, None)
        res["corr_simt_ms"] = med
    xc.set_impl(0)

    n = out.numel()
    med, best = timeit(lambda: ops.acc(acc, out, n, False), iters, flush)
    res["acc_ms"] = med
    res["acc_gbs"] = 12 * n / med / 1e6
    med, best = timeit(lambda: ops.acc(acc, out, n, True), iters, flush)
    res["acc_first_ms"] = med
    res["acc_first_gbs"] = 8 * n / med / 1e6

    nvis = 4704
    a2i = np.arange(ninput, dtype=np.int32).reshape(nstand, npol)
    a2bl, conj = ops.get_order(a2i)
    rng = np.random.default_rng(0)
    sel = rng.integers(0, a2bl.size, nvis)
    vm = torch.from_numpy(a2bl.reshape(-1)[sel].copy()).cuda()
    cj = torch.from_numpy(conj.reshape(-1)[sel].copy()).cuda()
    so = torch.zeros((nchan // 4, nvis, 2), dtype=torch.int32, device=dev)
    med, best = timeit(lambda: ops.subsel(out, so, vm, cj, nvis, nchan, 4, W), iters, flush)
    res["subsel_ms"] = med
    res["subsel_alg_gbs"] = (8 * nvis * nchan + 8 * nvis * nchan // 4) / med / 1e6

    if nchan <= 192:
        ro = torch.zeros((nstand, nstand, 2, 2, nchan, 2), dtype=torch.int32, device=dev)
        da, dc = torch.from_numpy(a2bl).cuda(), torch.from_numpy(conj).cuda()
        med, best = timeit(lambda: ops.reorder(out, ro, da, dc, nstand, npol, nchan, W), max(2, iters // 3), flush)
        res["reorder_ms"] = med
        res["reorder_gbs"] = (ro.numel() * 4 * 1.5) / med / 1e6
        del ro

    for nbeam, nt in ((32, 960), (16, 480)):
        bf = ops.Beamformer(ninput=ninput, nchan=nchan, ntime=nt, nbeam=nbeam)
        vv = torch.randint(0, 256, (nt, nchan, ninput), dtype=torch.uint8, device=dev)
        w = torch.randn((nchan, nbeam, ninput, 2), dtype=torch.float32, device=dev)
        bo = torch.zeros((nchan, nbeam, nt, 2), dtype=torch.float32, device=dev)
        po = torch.zeros((nbeam // 2, nt // 24, nchan, 4), dtype=torch.float32, device=dev)
        med, best = timeit(lambda: bf.run(vv, bo, w), iters, flush)
        tag = "bf%d_%d" % (nbeam, nt)
        res[tag + "_ms"] = med
        res[tag + "_gsps"] = nt * nchan * ninput / med / 1e6
        res[tag + "_tflops"] = 8.0 * nbeam * ninput * nchan * nt / med / 1e9
        med, best = timeit(lambda: bf.integrate(bo, po, 24), iters, flush)
        res[tag + "_integ_ms"] = med
        res[tag + "_integ_gbs"] = (bo.numel() * 4 + po.numel() * 4) / med / 1e6
        bf.close()
    print(json.dumps(res, indent=1))


if __name__ == "__main__":
    main()
