#!/usr/bin/env python
"""X-engine benchmark (contract: see the task statement / DESIGN.md "Measurement").

N = 1 (BASELINE.json configs[1]): Corr + CorrAcc on one B200 -- 704 inputs x 192 channels, 480-sample gulps,
accumulated to 250-gulp (~5 s of sky) dumps.  One step = one 480-sample gulp (64 880 640 4+4-bit complex samples)
through the fused correlate+accumulate kernel; every `--acc-gulps` steps the integration is dumped (device copy for
`value`, D2H to pinned memory for `e2e`).
N > 1 (configs[2] + configs[4]): every rank owns a 384-channel slab of the 3072-channel band (weak scaling, no
data-path collective).  One step = one 480-sample gulp through the fused correlate + accumulate + sub-select kernel;
EVERY step the rank's fast dump (int32 [96][4704][2], 3.6 MB) is pushed to rank 0 by peer writes over NVLink inside
the timed region (the exchange the reference makes as UDP packets every integration, corr_subsel_block.py:298-305);
every `--acc-gulps` steps the 766 MB accumulated visibilities are gathered to rank 0 too.
The same JSON line carries `xengine_full`: the whole X-engine (Corr+CorrAcc+CorrSubsel AND Beamform+power on the same
gulps, two streams; N > 1: fast dumps and beam powers pushed to rank 0 every step) against the 51.74 Gsamples/s
full-band real-time rate, and `blocks_graph`: the same work driven through the Python blocks on the ring shim.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
"""
import argparse
import json
import os
import subprocess
import struct
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "caltech-bifrost-dsp_b200"))
sys.path.insert(0, ROOT)

NSTAND, NPOL = 352, 2
NINPUT = NSTAND * NPOL
OPS_PER_SAMPLE = 8 * 248512 / NINPUT          # 2824 int ops per input sample
INT8_NOMINAL_TOPS = 4500.0                    # B200 dense int8, datasheet
METRIC = "xengine_input_gsamples_per_s"
UNIT = "Gsamples/s"


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=250)
    p.add_argument("--warmup", type=int, default=10)
    p.add_argument("--impl", default="b200", choices=["b200", "reference"])
    p.add_argument("--nchan-per-gpu", type=int, default=None,
                   help="default: 192 at --gpus 1 (configs[1]), 384 at --gpus > 1 (3072 / 8, configs[2])")
    p.add_argument("--ntime", type=int, default=480)
    p.add_argument("--acc-gulps", type=int, default=250)
    p.add_argument("--nccl-gather", action="store_true",
                   help="gather dumps with dist.gather (NCCL kernels) instead of copy-engine peer writes")
    p.add_argument("--push", default="sm", choices=["sm", "ce"],
                   help="per-step products to rank 0: xeng_peer_push (SM stores) or cudaMemcpyAsync (copy engines)")
    p.add_argument("--push-slots", type=int, default=4,
                   help="N > 1: fast-dump buffers a rank may have in flight to rank 0 before its correlator waits")
    p.add_argument("--dump-slices", type=int, default=None,
                   help="N > 1: the long dump leaves for rank 0 in this many per-step slices of the next integration "
                        "(default acc_gulps // 4; 0 = one burst)")
    p.add_argument("--cpu-seconds", type=float, default=15.0, help="budget for the cpu_baseline sample")
    p.add_argument("--no-cpu-baseline", action="store_true")
    p.add_argument("--no-path-kernels", action="store_true", help="skip the per-kernel figures of the rest of the path")
    p.add_argument("--no-xengine-full", action="store_true", help="skip the whole-X-engine and block-graph sections")
    p.add_argument("--full-steps", type=int, default=40, help="steps of the whole-X-engine section")
    a = p.parse_args()
    if a.nchan_per_gpu is None:
        a.nchan_per_gpu = 192 if a.gpus == 1 else 384
    return a


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            d = json.load(f)
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


# ----------------------------------------------------------------------------------------------
# CPU arm: the reference's own CPU formulation of the path (numpy; bifrost/xGPU are absent submodules)
# ----------------------------------------------------------------------------------------------
def cpu_corr_sample(nchan_total, ntime, budget_s, seed=0xdeadbeef, wrap=False, gulp=None):
    """Time oracle Corr (xGPU order) + CorrAcc on a bounded sample of the workload: a slab of channels of
    one gulp, grown until ~budget_s of CPU work has been spent (wrap=True carries on into further gulps when
    the host finishes one gulp inside the budget).  `gulp`: uint8 [ntime][nchan][704], the GPU leg's own first gulp
    ("same synthetic input"); without it (reference arm: no GPU in the process) bytes are drawn as the
    reference's generator draws them (make_golden_inputs.py:137).  Returns (Gsamples/s, description, cores, s, samples)."""
    from oracle import xeng_oracle as orc
    rng = np.random.RandomState(seed & 0xFFFFFFFF)
    nch = 2

    def slab(c):
        if gulp is None:
            return rng.randint(0, 255, size=(ntime, nch, NSTAND, NPOL), dtype=np.uint8)
        c %= gulp.shape[1] - nch + 1
        return np.ascontiguousarray(gulp[:, c:c + nch]).reshape(ntime, nch, NSTAND, NPOL)

    orc.corr_xgpu(slab(0))                                                     # warm-up (BLAS threads, index maps)
    done_ch, t_used = 0, 0.0
    acc = None
    while t_used < budget_s and (wrap or done_ch < nchan_total):
        v = slab(done_ch)
        t0 = time.perf_counter()
        x = orc.corr_xgpu(v)
        acc = orc.corr_acc(x.copy() if acc is None else acc, x, first=acc is None)
        t_used += time.perf_counter() - t0
        done_ch += nch
    samples = done_ch * ntime * NINPUT
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    desc = "%d channels = %.2f gulps of %d channels x %d samples, 704 inputs (numpy/BLAS float32-exact matmul " \
           "port of Corr._test + CorrAcc, all host threads%s), %.1f s" % (
               done_ch, done_ch / nchan_total, nchan_total, ntime,
               ", the GPU leg's own voltages" if gulp is not None else "", t_used)
    return samples / t_used / 1e9, desc, cores, t_used, samples


def cpu_literal_sample(gulp, nt_sub=96):
    """The reference's literal verification loop (Corr._test, corr_block.py:279-284: four np.outer per (time, channel))
    on a stated subset -- one channel x nt_sub samples of the GPU leg's first gulp -- on one core; checked equal to
    the BLAS port on the same bytes."""
    from oracle import xeng_oracle as orc
    v = np.ascontiguousarray(gulp[:nt_sub, :1]).reshape(nt_sub, 1, NSTAND, NPOL)
    t0 = time.perf_counter()
    lit = orc.corr_loop_literal(v)
    dt = time.perf_counter() - t0
    fre, fim = orc.corr_full(v)                       # [chan, s0, s1, p0, p1] = sum v(s0,p0) conj(v(s1,p1))
    same = True
    for p0 in range(NPOL):
        for p1 in range(NPOL):                        # literal: out[c, i, j, 2 p0 + p1] = sum conj(v(i,p0)) v(j,p1)
            want = fre[:, :, :, p1, p0].transpose(0, 2, 1) + 1j * fim[:, :, :, p1, p0].transpose(0, 2, 1)
            same = same and bool(np.array_equal(lit[:, :, :, 2 * p0 + p1], want))
    return {"value": nt_sub * NINPUT / dt / 1e9, "unit": UNIT, "cores": 1, "kind": "port",
            "sample": "literal np.outer loop of corr_block.py:279-284, 1 channel x %d samples x 704 inputs, %.2f s "
                      "(a 480 x 192 gulp at this rate: %.0f s)" % (nt_sub, dt, dt / nt_sub * 480 * 192),
            "equals_blas_port": same}


def verify_accumulator(orc, d_acc, nchan, wpc, counts, gulps_host, chans):
    """The timed accumulator against the oracle on the SAME bytes: channels `chans` of sum_k counts[k] X(gulp k)."""
    import torch
    got = d_acc.view(2, nchan, wpc)[:, chans].cpu().numpy()
    want = np.zeros(got.shape, dtype=np.int64)
    for k, n in enumerate(counts):
        if n:
            v = np.ascontiguousarray(gulps_host[k][:, chans]).reshape(-1, len(chans), NSTAND, NPOL)
            want += int(n) * orc.corr_xgpu(v).astype(np.int64)
    return bool(np.array_equal(got, want.astype(np.int32)))


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    nchan = args.nchan_per_gpu * args.gpus
    per_step_budget = max(1.0, min(args.cpu_seconds, 120.0 / max(1, args.steps + args.warmup)))
    for _ in range(min(args.warmup, 1)):
        cpu_corr_sample(nchan, args.ntime, per_step_budget)
    tot_t, tot_s, desc, cores = 0.0, 0, "", 1
    for _ in range(args.steps):
        _, desc, cores, t, s = cpu_corr_sample(nchan, args.ntime, per_step_budget)
        tot_t += t
        tot_s += s
    val = tot_s / tot_t / 1e9
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * tot_t / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int8->int32",
        "data": "synthetic",
        "config": workload_config(args),
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port", "sample": "per step: " + desc},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


def workload_config(args):
    return {
        "workload": "Corr+CorrAcc%s: 704 inputs (352 stands x 2 pol) x %d channels/GPU, %d-sample gulps of 4+4-bit "
                    "voltages, fused correlate+accumulate%s, dump every %d gulps" % (
                        "+CorrSubsel (BASELINE configs[2]+[4])" if args.gpus > 1 else " (BASELINE configs[1])",
                        args.nchan_per_gpu, args.ntime,
                        "+sub-select with the fast dump shipped to rank 0 every step" if args.gpus > 1 else "",
                        args.acc_gulps),
        "nchan_per_gpu": args.nchan_per_gpu, "ninput": NINPUT, "ntime_gulp": args.ntime,
        "acc_gulps": args.acc_gulps,
        "sharding": ("channel slabs of the 3072-channel band, 1 rank per GPU over torch.distributed/NCCL; EVERY step each "
                     "rank's fast dump (CorrSubsel's 4704 baselines, fused into the correlator launch) is pushed to "
                     "rank 0 over NVLink inside the timed region; each long dump is gathered to rank 0 on a side stream")
                    if args.gpus > 1 else "single GPU",
        "l2": "inputs rotate over 4 distinct gulps (>=260 MB) and every step read-modify-writes the 383 MB "
              "accumulator, both larger than the 126 MB L2",
    }


# ----------------------------------------------------------------------------------------------
# clocks sampling (B200_PROFILING.md recipe)
# ----------------------------------------------------------------------------------------------
class ClockSampler(object):
    Q = "index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active," \
        "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q,
                                       "--format=csv,noheader,nounits", "-lms", "100"],
                                      stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            self.p = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.f.read().splitlines():
            c = [x.strip() for x in ln.split(",")]
            if len(c) < 9:
                continue
            try:
                sm.append(float(c[1]))
                mx.append(float(c[2]))
            except ValueError:
                continue
            for n, val in zip(names, c[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(n)
        self.f.close()
        os.unlink(self.f.name)
        if sm:
            # under load = samples in the upper half of what was seen (idle samples bracket the run)
            out["sm_mhz"] = float(np.median(sm))
            out["sm_max_mhz"] = float(max(mx))
            out["samples"] = len(sm)
        out["reasons"] = sorted(reasons)
        return out


# ----------------------------------------------------------------------------------------------
# the other kernels of the path: bandwidth / flop figures next to the headline kernel
# ----------------------------------------------------------------------------------------------
def ncu_traffic(nchan, ntime):
    """dram__bytes_read.sum + dram__bytes_write.sum of one fused Corr+CorrAcc launch, from the committed ncu
    --set full capture (profiles/corr_traffic.json, written by scripts/summarize_ncu.py)."""
    path = os.path.join(ROOT, "profiles", "corr_traffic.json")
    if os.path.exists(path):
        with open(path) as f:
            d = json.load(f)
        if d.get("nchan") == nchan and d.get("ntime_gulp") == ntime:     # captured on this workload only
            return float(d["dram_bytes_per_launch"]), d.get("source")
    return None, None


def path_kernels(torch, ops, dev, nchan, hbm_peak, iters=10):
    """CUDA-event timings (L2 flushed between launches) of CorrAcc, CorrSubsel, reorder, Beamform and BeamformSum at
    the sizes of BASELINE configs[1]/[3]/[4]; algorithmic bytes / flops per SURVEY 8(d)."""
    flush = torch.zeros(256 << 20, dtype=torch.uint8, device=dev)

    def timeit(fn, n=iters):
        for _ in range(2):
            fn()
        torch.cuda.synchronize(dev)
        ts = []
        for _ in range(n):
            flush.add_(1)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize(dev)
            ts.append(e0.elapsed_time(e1))
        return float(np.median(ts))

    res = {}
    wpc = ops.words_per_chan(NSTAND)
    nw = 2 * nchan * wpc
    vis = torch.randint(-1000, 1000, (nw,), dtype=torch.int32, device=dev)
    acc = torch.zeros(nw, dtype=torch.int32, device=dev)
    ms = timeit(lambda: ops.acc(acc, vis, nw, False))
    res["corr_acc"] = {"ms": ms, "alg_bytes": 12 * nw, "gbs": 12 * nw / ms / 1e6, "frac_hbm": 12 * nw / ms / 1e6 / hbm_peak}
    a2i = np.arange(NINPUT, dtype=np.int32).reshape(NSTAND, NPOL)
    a2bl, conj = ops.get_order(a2i)
    from oracle import xeng_oracle as orc_
    real_vm, real_cj = real_selection(ops, orc_, np.random.default_rng(4704))
    for nvis in (4704, 4656):
        if nvis == 4704:      # CorrSubsel's shape: 48 stands, all pairs, 4 pol products each = whole 16-byte cells
            vm, cj = torch.from_numpy(real_vm).to(dev), torch.from_numpy(real_cj).to(dev)
        else:                 # the docs' count as random single words: every word in a cell of its own (worst case)
            sel = np.random.default_rng(nvis).integers(0, a2bl.size, nvis)
            vm = torch.from_numpy(a2bl.reshape(-1)[sel].copy()).to(dev)
            cj = torch.from_numpy(conj.reshape(-1)[sel].copy()).to(dev)
        so = torch.zeros((nchan // 4, nvis, 2), dtype=torch.int32, device=dev)
        ms = timeit(lambda: ops.subsel(vis, so, vm, cj, nvis, nchan, 4, wpc))
        ab = 8 * nvis * nchan + 8 * nvis * nchan // 4
        res["corr_subsel_%d" % nvis] = {"ms": ms, "alg_bytes": ab, "gbs": ab / ms / 1e6,
                                        "selection": "48 stands x all pairs x 4 pol products (whole cells)" if nvis == 4704
                                        else "random single words",
                                        "sector_bytes": 2 * 32 * (nvis // 4 if nvis == 4704 else nvis) * nchan + 8 * nvis * nchan // 4}
    ro = torch.zeros((NSTAND, NSTAND, 2, 2, nchan, 2), dtype=torch.int32, device=dev)
    da, dc = torch.from_numpy(a2bl).to(dev), torch.from_numpy(conj).to(dev)
    ms = timeit(lambda: ops.reorder(vis, ro, da, dc, NSTAND, NPOL, nchan, wpc), 4)
    ab = ro.numel() * 4 + ro.numel() * 4 // 2     # every output word written, each stored word read once
    res["reorder"] = {"ms": ms, "alg_bytes": ab, "gbs": ab / ms / 1e6, "frac_hbm": ab / ms / 1e6 / hbm_peak}
    nbl = NSTAND * (NSTAND + 1) // 2
    pk = torch.zeros((nbl, nchan, 2, 2, 2), dtype=torch.int32, device=dev)
    ms = timeit(lambda: ops.reorder_packets(vis, pk, da, dc, NSTAND, NPOL, nchan, wpc, ops.PAYLOAD_COR), 4)
    ab = pk.numel() * 4 * 2                       # each valid word read once, written once
    res["reorder_cor_payloads"] = {"ms": ms, "alg_bytes": ab, "gbs": ab / ms / 1e6, "frac_hbm": ab / ms / 1e6 / hbm_peak}
    del ro, pk, vis, acc
    # fused Corr + CorrAcc + CorrSubsel step (configs[4]): one launch per gulp, 4704 baselines summed over 4 channels
    ntime = 480
    xc = ops.XCorr(nstand=NSTAND, npol=NPOL, nchan=nchan, ntime_gulp=ntime)
    gulps = [torch.randint(0, 256, (ntime, nchan, NINPUT), dtype=torch.uint8, device=dev) for _ in range(4)]
    acc2 = torch.zeros(nw, dtype=torch.int32, device=dev)
    xc.set_subsel(real_vm, real_cj, 4)
    fast = torch.zeros((nchan // 4, 4704, 2), dtype=torch.int32, device=dev)
    k = [0]

    def fused():
        xc.accumulate_subsel(gulps[k[0] % 4], acc2, False, fast)
        k[0] += 1
    def plain():
        xc.accumulate(gulps[k[0] % 4], acc2, False)
        k[0] += 1
    ms_plain = timeit(plain, 20)
    ms = timeit(fused, 20)
    ab = ntime * nchan * NINPUT + 2 * 4 * nw + 8 * 4704 * nchan // 4
    res["corr_acc_subsel_fused"] = {"ms": ms, "gsamples_per_s": ntime * nchan * NINPUT / ms / 1e6, "alg_bytes": ab,
                                    "gbs": ab / ms / 1e6, "frac_hbm": ab / ms / 1e6 / hbm_peak,
                                    "corr_acc_only_ms": ms_plain, "unfused_ms": None}
    # the same work as the three separate blocks do it: Corr (writes the per-gulp matrix) + CorrAcc + CorrSubsel
    ms_store = timeit(lambda: xc.run(gulps[0], acc2, True), 10)
    res["corr_store"] = {"ms": ms_store, "gsamples_per_s": ntime * nchan * NINPUT / ms_store / 1e6}
    res["corr_acc_subsel_fused"]["unfused_ms"] = ms_store + res["corr_acc"]["ms"] + res["corr_subsel_4704"]["ms"]
    xc.close()
    # a span of five gulps in ONE launch (ntime 2400, Corr's constructor default acc_len, corr_block.py:221): one
    # accumulator read-modify-write per 2400 samples instead of five -- the tensor-bound regime of the same kernel
    nt5 = 5 * ntime
    span = torch.randint(0, 256, (nt5, nchan, NINPUT), dtype=torch.uint8, device=dev)
    xc5 = ops.XCorr(nstand=NSTAND, npol=NPOL, nchan=nchan, ntime_gulp=nt5)
    ms5 = timeit(lambda: xc5.accumulate(span, acc2, False), 10)
    ops5 = 8.0 * 248512 * nchan * nt5
    res["corr_acc_span_2400"] = {"ms": ms5, "gsamples_per_s": nt5 * nchan * NINPUT / ms5 / 1e6,
                                 "useful_int8_tops": ops5 / ms5 / 1e9, "frac_int8_nominal": ops5 / ms5 / 1e9 / INT8_NOMINAL_TOPS,
                                 "note": "five 480-sample gulps per launch: not the headline's configuration"}
    xc5.close()
    del gulps, acc2, fast, span
    # ingest: one gulp of SNAP2 packets (96 ch x 64 inputs each, 22 per spectrum) -> voltage buffer
    ncp, npp, stride = 96, 64, 32 + 96 * 64
    npk = ntime * (nchan // ncp) * (NINPUT // npp)
    hdr = np.zeros((npk, 32), dtype=np.uint8)
    k = 0
    for t in range(ntime):
        for cb in range(nchan // ncp):
            for pb in range(NINPUT // npp):
                hdr[k] = np.frombuffer(struct.pack('>QLHHHHLLL', 7000 + t, 0, npp, NINPUT, ncp, nchan, cb, cb * ncp, pb * npp),
                                       dtype=np.uint8)
                k += 1
    raw = torch.randint(0, 256, (npk, stride), dtype=torch.uint8, device=dev)
    raw[:, :32] = torch.from_numpy(hdr).to(dev)
    vbuf = torch.zeros((ntime, nchan, NINPUT), dtype=torch.uint8, device=dev)
    ms = timeit(lambda: ops.snap2_assemble(raw, npk, stride, vbuf, 7000, ntime, nchan, NINPUT, 0, False))
    ab = 2 * ntime * nchan * NINPUT
    res["snap2_assemble"] = {"ms": ms, "packets": npk, "alg_bytes": ab, "gbs": ab / ms / 1e6,
                             "frac_hbm": ab / ms / 1e6 / hbm_peak, "gsamples_per_s": ntime * nchan * NINPUT / ms / 1e6}
    del raw, vbuf
    nbeam, nt, nsum = 32, 960, 24
    bf = ops.Beamformer(ninput=NINPUT, nchan=nchan, ntime=nt, nbeam=nbeam)
    vv = torch.randint(0, 256, (nt, nchan, NINPUT), dtype=torch.uint8, device=dev)
    w = torch.randn((nchan, nbeam, NINPUT, 2), dtype=torch.float32, device=dev)
    bo = torch.zeros((nchan, nbeam, nt, 2), dtype=torch.float32, device=dev)
    po = torch.zeros((nbeam // 2, nt // nsum, nchan, 4), dtype=torch.float32, device=dev)
    ms = timeit(lambda: bf.run(vv, bo, w))
    fl = 8.0 * nbeam * NINPUT * nchan * nt
    res["beamform"] = {"ms": ms, "gsamples_per_s": nt * nchan * NINPUT / ms / 1e6, "tflops_fp32": fl / ms / 1e9,
                       "config": "%d beams x %d inputs x %d ch x %d samples" % (nbeam, NINPUT, nchan, nt)}
    ms = timeit(lambda: bf.integrate(bo, po, nsum))
    ab = bo.numel() * 4 + po.numel() * 4
    res["beamform_sum"] = {"ms": ms, "alg_bytes": ab, "gbs": ab / ms / 1e6, "frac_hbm": ab / ms / 1e6 / hbm_peak}
    # configs[3] as one kernel: power integration in the beamformer's epilogue, voltage beams never written
    ms_f = timeit(lambda: bf.run_integrate(vv, po, w, nsum))
    res["beamform_power_fused"] = {"ms": ms_f, "gsamples_per_s": nt * nchan * NINPUT / ms_f / 1e6,
                                   "tflops_fp32": fl / ms_f / 1e9, "unfused_ms": res["beamform"]["ms"] + ms}
    res["beamform"]["note"] = "tcgen05 int8 digit-plane kernel, weights converted on EVERY call (bfBeamformRun's convention)"
    bf.set_weights(w)            # converted once, as the Beamform block does at each coefficient update
    ms_p = timeit(lambda: bf.run(vv, bo, None))
    ms_pf = timeit(lambda: bf.run_integrate(vv, po, None, nsum))
    ms_c = timeit(lambda: bf.set_weights(w))
    alg = vv.numel() + bo.numel() * 4 + 6 * nbeam * 2 * NINPUT * nchan
    res["beamform_weights_preloaded"] = {"ms": ms_p, "gsamples_per_s": nt * nchan * NINPUT / ms_p / 1e6,
                                         "fp32_equiv_tflops": fl / ms_p / 1e9, "int8_tops": 3 * fl / ms_p / 1e9,
                                         "alg_bytes": alg, "gbs": alg / ms_p / 1e6, "frac_hbm": alg / ms_p / 1e6 / hbm_peak,
                                         "power_fused_ms": ms_pf, "weight_conversion_ms": ms_c}
    bf.set_impl(1)
    ms_32 = timeit(lambda: bf.run(vv, bo, w), 4)
    res["beamform_fp32_kernel"] = {"ms": ms_32, "tflops_fp32": fl / ms_32 / 1e9,
                                   "note": "the CUDA-core cross-check implementation (r1's kernel)"}
    bf.close()
    del vv, w, bo, po
    # offline upchannelise path at the scripts' sizes: 480-sample gulp = 15 frames x 32 fine samples,
    # 96 ch (lwa352-upchan-bf.py:86-87, --nchan default), 4 beams (--nbeam default)
    nfr, nf, nch_u, nb_u = 15, 32, 96, 4
    vu = torch.randint(0, 256, (nfr, nf, nch_u, NINPUT), dtype=torch.uint8, device=dev)
    xu = torch.zeros((nfr, nch_u, NINPUT, nf, 2), dtype=torch.float32, device=dev)
    wu = torch.randn((nb_u, nch_u, NINPUT, nf, 2), dtype=torch.float32, device=dev)
    ou = torch.zeros((nfr, nb_u, nch_u, nf, 2), dtype=torch.float32, device=dev)
    nsamp = nfr * nf * nch_u * NINPUT
    for name, lay in (("upchan_fft_fine_last", ops.UPCHAN_FINE_LAST), ("upchan_fft_freq_merged", ops.UPCHAN_FREQ_MERGED)):
        ms = timeit(lambda: ops.upchan_fft(vu, xu, nfr, nf, nch_u, NINPUT, lay))
        ab = 9 * nsamp
        res[name] = {"ms": ms, "alg_bytes": ab, "gbs": ab / ms / 1e6, "frac_hbm": ab / ms / 1e6 / hbm_peak,
                     "gsamples_per_s": nsamp / ms / 1e6}
    ms = timeit(lambda: ops.upchan_beamform(vu, wu, ou, nfr, nf, nch_u, NINPUT, nb_u))
    ab = nsamp + wu.numel() * 4 + ou.numel() * 4
    res["upchan_beamform_fused"] = {"ms": ms, "alg_bytes": ab, "gbs": ab / ms / 1e6,
                                    "frac_hbm": ab / ms / 1e6 / hbm_peak, "gsamples_per_s": nsamp / ms / 1e6,
                                    "config": "%d frames x %d fine x %d ch x %d inputs, %d beams" % (nfr, nf, nch_u, NINPUT, nb_u)}
    del vu, xu, wu, ou
    # the imaging script's correlate (lwa352-upchan-imag.py:85-106): 480 samples at NUPCHAN = 2 = 240 frames, the first
    # 192 merged frequencies, complex fp32 Hermitian update on the FMA pipe
    nfr_c, nfq_c = 240, 192
    xc_in = torch.randn((nfr_c, nfq_c, NINPUT, 2), dtype=torch.float32, device=dev)
    xc_out = torch.zeros((nfq_c, NINPUT, NINPUT, 2), dtype=torch.float32, device=dev)
    ms = timeit(lambda: ops.upchan_correlate(xc_in, xc_out, nfr_c, nfq_c, NINPUT, False), 4)
    T_ = (NINPUT + 63) // 64
    fl = 8.0 * nfr_c * nfq_c * (T_ * (T_ + 1) // 2) * 64 * 64
    res["upchan_correlate"] = {"ms": ms, "tflops_fp32": fl / ms / 1e9, "gsamples_per_s": nfr_c * nfq_c * NINPUT / ms / 1e6,
                               "config": "%d frames x %d freqs x %d inputs, accumulating" % (nfr_c, nfq_c, NINPUT)}
    return res


# ----------------------------------------------------------------------------------------------
# B200 arm
# ----------------------------------------------------------------------------------------------
def stats(x):
    x = np.asarray(x, dtype=np.float64)
    return {"min": float(x.min()), "median": float(np.median(x)), "max": float(x.max())}


def real_selection(ops, orc, rng, nvis_stands=48):
    """CorrSubsel's selection shape (corr_subsel_block.py:185: 48 x 49 / 2 stand pairs x 4 pol products = 4704) for
    48 random stands: whole 16-byte cells, both station orders."""
    a2bl, conj = ops.get_order(np.arange(NINPUT, dtype=np.int32).reshape(NSTAND, NPOL))
    st = rng.choice(NSTAND, nvis_stands, replace=False)
    bls = [[[int(s0), p0], [int(s1), p1]] for i, s0 in enumerate(st) for s1 in st[:i + 1]
           for p0 in range(2) for p1 in range(2)]
    vm, cj = orc.baselines_to_vismap(bls, a2bl, conj)
    return vm, cj


def run_b200(args):
    import torch
    import torch.distributed as dist
    from xengine import ops, sharding
    from oracle import xeng_oracle as orc          # the checker: verification of what was timed + the CPU leg

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the B200 arm has no CPU fallback (use --impl reference)")
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    real_stdout = sys.stdout
    if world > 1:
        # NCCL / c10d print a version banner on fd 1: keep stdout to the one JSON line by pointing fd 1 at stderr
        # for the lifetime of the process group and writing the result to a duplicate of the original stdout.
        sys.stdout.flush()
        real_stdout = os.fdopen(os.dup(1), "w")
        os.dup2(2, 1)
        dist.init_process_group("nccl", device_id=dev)
    nchan, ntime, K, Wm, accg = args.nchan_per_gpu, args.ntime, args.steps, args.warmup, args.acc_gulps
    wpc = ops.words_per_chan(NSTAND)
    nsamp_step = ntime * nchan * NINPUT                      # per GPU
    in_bytes = nsamp_step
    out_words = 2 * nchan * wpc
    fused_fast = world > 1                                   # N > 1: the step also produces + ships the fast dump
    NVIS, NSUM = 4704, 4

    # synthetic voltages: 4 distinct gulps, uniform bytes, fixed seed (per rank = per channel slab)
    g = torch.Generator(device=dev)
    g.manual_seed(0xdeadbeef + rank)
    NBUF = 4
    d_in = [torch.randint(0, 256, (ntime, nchan, NINPUT), dtype=torch.uint8, device=dev, generator=g)
            for _ in range(NBUF)]
    gulps_host = [x.cpu().numpy() for x in d_in]
    # CorrAcc's output ring: as in blocks/corr_block.py (acc_oring=) the fused launch accumulates straight into the
    # span it will commit, so a finished integration is shipped from where it lies while the next one runs in the
    # other span -- no snapshot copy on the correlator's stream
    accs = [torch.zeros(out_words, dtype=torch.int32, device=dev) for _ in range(2)]
    cur = [0]

    def acc():
        return accs[cur[0]]

    d_dump = torch.zeros(out_words, dtype=torch.int32, device=dev) if world == 1 else None   # N = 1: the dump's destination
    dumped = [None]
    gather_list = None
    stream = torch.cuda.Stream(device=dev)
    xc = ops.XCorr(nstand=NSTAND, npol=NPOL, nchan=nchan, ntime_gulp=ntime, gpu=local)
    xc.set_stream(stream)
    vm, cj = real_selection(ops, orc, np.random.default_rng(4704))
    xc.set_subsel(vm, cj, NSUM)
    NSLOT = max(2, args.push_slots)
    fast = [torch.zeros((nchan // NSUM, NVIS, 2), dtype=torch.int32, device=dev) for _ in range(NSLOT)]

    def barrier():
        # synchronize on BOTH sides of the rendezvous: a rank's side-stream traffic (a 766 MB dump still streaming
        # into rank 0) must have landed before any rank starts the next timed region, or the early ranks' first
        # pushes queue behind it (r02o: one 0.6 ms step on every non-root rank)
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    # ---- exchanges (N > 1): long dump -> PeerGather, per-step fast dump -> PeerStream; NCCL fallbacks
    gstream = torch.cuda.Stream(device=dev)                              # long dumps
    pstream = torch.cuda.Stream(device=dev) if world > 1 else None      # per-step pushes
    peer = fpush = None
    push_kind = None
    if world > 1:
        if not args.nccl_gather:
            try:
                peer = sharding.PeerGather(out_words, torch.int32, dev, dist, root=0)
                fpush = sharding.PeerStream(fast[0].numel(), torch.int32, dev, dist, nslot=NSLOT, root=0, via=args.push)
            except Exception as e:                              # no symmetric memory on this node: all ranks use NCCL
                sys.stderr.write("rank %d: symmetric-memory peer writes unavailable (%s)\n" % (rank, e))
                peer = fpush = None
        agreed = torch.tensor([1 if (peer is not None and fpush is not None) else 0], device=dev)
        dist.all_reduce(agreed, op=dist.ReduceOp.MIN)
        if int(agreed) == 0:
            peer, args.nccl_gather = None, True
            fpush = sharding.GatherStream(fast[0].numel(), torch.int32, dev, dist, nslot=NSLOT, root=0)
            push_kind = "NCCL dist.gather every step"
        else:
            push_kind = ("peer writes into rank 0's symmetric-memory slots over NVLink (%s), no cross-rank wait"
                         % ("xeng_peer_move: SM stores from CTAs resident beside the correlator, the buffer left zeroed for "
                            "the launch that adds into it next" if args.push == "sm"
                            else "cudaMemcpyAsync: copy engines"))
    if world > 1 and rank == 0 and args.nccl_gather:
        gather_list = [torch.empty_like(accs[0]) for _ in range(world)]

    def dump_device():
        # The finished integration leaves from its span on a side stream -- N = 1: CorrAcc's copy_array into the
        # consumer's buffer; N > 1: the gather to rank 0 -- while the correlator carries on with the next integration
        # in the other span, as in the pipeline.  The next dump waits for this one (so a span is never reused while
        # it travels), and so does the end of the timed region.
        src = acc()
        flush_dump()
        stream.wait_stream(gstream)
        dumped[0] = src
        if peer is not None and nslices > 0:
            # in slices, one per step of the next integration (dump_slice): 766 MB per rank in one burst fills the
            # links into rank 0 for 7 ms at 8 ranks and the per-step products wait behind it (profiles/r02x)
            pending["src"], pending["ranges"] = src, peer.slices(nslices)
            dseq[0] += 1
            return
        gstream.wait_stream(stream)
        with torch.cuda.stream(gstream):
            if world == 1:
                d_dump.copy_(src, non_blocking=True)
            elif peer is not None:
                peer.push(src)
            else:
                sharding.gather_to_root(src, dist, root=0, out=gather_list)

    nslices = (accg // 4 if args.dump_slices is None else args.dump_slices) if peer is not None else 0
    pending = {"src": None, "ranges": []}
    dseq = [0]                                               # sliced dumps sent so far = the sequence word of the last

    def dump_slice(after):
        """The next slice of the travelling dump, on the dump stream, once the step that recorded `after` is done."""
        lo, hi = pending["ranges"].pop(0)
        with torch.cuda.stream(gstream):
            gstream.wait_event(after)
            peer.push_range(pending["src"], lo, hi)
            if not pending["ranges"]:
                peer.commit(dseq[0])                           # a sequence word behind the last slice: no rank waits
                pending["src"] = None

    def flush_dump():
        """Whatever is left of a travelling dump leaves now (end of a timed region; a dump timed on its own)."""
        if pending["src"] is not None:
            ev = torch.cuda.Event()
            ev.record(stream)
            while pending["src"] is not None:
                dump_slice(ev)

    push_launches = 2 if getattr(fpush, "via", None) == "sm" else 0
    push_moves = push_launches > 0           # SM pushes move the fast dump: no per-step memset before the correlator
    kept = [torch.zeros_like(fast[0]) for _ in range(2)] if fused_fast else [None, None]   # [timed region, overlap runs]
    fast_done = [torch.cuda.Event() for _ in range(NSLOT)]
    fast_free = [torch.cuda.Event() for _ in range(NSLOT)]
    counts = [0] * NBUF                                      # gulps of each buffer in the running integration
    nlaunch = [0]
    seqc = [0]

    def launch_step(src, k_buf, first, keep=None):
        """One step on `stream` (current): the hot-path launch (+ N > 1: the fast dump's push on the side stream).
        Returns the step's sequence number.  `keep`: a buffer that receives a copy of this step's fast dump (the dump
        itself is MOVED to rank 0: xeng_peer_move leaves the buffer zeroed for the launch that adds into it next)."""
        seq = seqc[0]
        seqc[0] += 1
        if first:                                            # a new integration opens in the other span
            cur[0] ^= 1
            for k in range(NBUF):
                counts[k] = 0
        counts[k_buf] += 1
        if not fused_fast:
            xc.accumulate(src, acc(), first)
            nlaunch[0] += 1
            return seq
        slot = seq % NSLOT
        stream.wait_event(fast_free[slot])                   # the push of step seq - NSLOT has left this buffer
        xc.accumulate_subsel(src, acc(), first, fast[slot], not push_moves)
        nlaunch[0] += 1
        if keep is not None:
            keep.copy_(fast[slot], non_blocking=True)
        fast_done[slot].record(stream)
        with torch.cuda.stream(pstream):
            pstream.wait_event(fast_done[slot])
            if push_moves:
                fpush.push(fast[slot], seq, move=True)
            else:
                fpush.push(fast[slot], seq)
            nlaunch[0] += push_launches                      # xeng_peer_push: payload kernel + sequence-word kernel
            fast_free[slot].record(pstream)
        if pending["src"] is not None:
            dump_slice(fast_done[slot])
        return seq

    # ---- value: inputs resident in HBM
    with torch.cuda.stream(stream):
        for i in range(Wm):
            launch_step(d_in[i % NBUF], i % NBUF, (i % accg) == 0)
        dump_device()          # warm the dump path too (first NCCL gather sets up its peer connections)
        flush_dump()
    barrier()
    sampler = ClockSampler(local) if rank == 0 else None
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(K + 1)]
    ndump = 0
    phase = accg // 2
    last_seq = -1
    with torch.cuda.stream(stream):
        # the timed region starts in the middle of a long integration (steady state): a dump falls every `accg`
        # steps, the first after accg - accg // 2, and travels to rank 0 while the next integration runs.
        # One UNTIMED step first: it opens a fresh integration (so the accumulator the timed region leaves behind is
        # a known sum, checked against the oracle below) and it is still running when the start event is recorded, so
        # the K timed launches queue up behind it back to back -- the first launch after a barrier otherwise shows the
        # host's wake-up latency, not the kernel (r01: +13 % on the max-over-ranks of a 20-step mean)
        launch_step(d_in[NBUF - 1], NBUF - 1, True)
        evs[0].record(stream)
        for i in range(K):
            g_i = i + phase
            last_seq = launch_step(d_in[i % NBUF], i % NBUF, (g_i % accg) == 0, kept[0] if i == K - 1 else None)
            evs[i + 1].record(stream)
            if (g_i % accg) == accg - 1:
                dump_device()
                ndump += 1
        flush_dump()
        stream.wait_stream(gstream)
        if world > 1:
            stream.wait_stream(pstream)
        end = torch.cuda.Event(enable_timing=True)
        end.record(stream)
    barrier()
    total_ms = evs[0].elapsed_time(end)
    step_ms = [evs[i].elapsed_time(evs[i + 1]) for i in range(K)]
    # kernel-only time: intervals that contain exactly one hot-path launch (those after a dump also contain the copy)
    kern = [step_ms[i] for i in range(K) if not (i > 0 and (((i - 1 + phase) % accg) == accg - 1))]
    kern_ms = float(np.median(kern)) if kern else total_ms / K
    xc.sync()
    # what was timed is what the reference computes: two channel slabs of the accumulator the timed region left
    # behind against the oracle on the same bytes
    vchans = [0, nchan // 2 - 1] if nchan > 2 else [0]
    verified = verify_accumulator(orc, acc(), nchan, wpc, counts, gulps_host, vchans)
    fast_verified = None
    if fused_fast:                                           # and the last fast dump, every channel group
        last = last_seq
        x_last = None
        sub_ch = [0, 1, 2, 3]                                # first output channel = input channels 0..3
        v4 = np.ascontiguousarray(gulps_host[(K - 1) % NBUF][:, sub_ch]).reshape(ntime, 4, NSTAND, NPOL)
        x_last = orc.corr_xgpu(v4)
        fast_verified = bool(np.array_equal(kept[0][0].cpu().numpy(), orc.subselect(x_last, vm, cj, NSUM)[0]))
    if not verified or fast_verified is False:
        raise RuntimeError("timed accumulator / fast dump differ from the oracle on the same voltages")

    # one dump on its own (copy + gather, nothing to overlap with): what each long integration adds at most
    barrier()
    with torch.cuda.stream(stream):
        dmp0 = torch.cuda.Event(enable_timing=True)
        dmp1 = torch.cuda.Event(enable_timing=True)
        dmp0.record(stream)
        dump_device()
        flush_dump()
        stream.wait_stream(gstream)
        dmp1.record(stream)
    barrier()
    dump_ms = dmp0.elapsed_time(dmp1)
    # what a dump costs when it travels beside the correlator, as it does at its true cadence: M steps with one dump
    # launched on the side stream at their start, against the same M steps alone (M steps outlast the dump)
    M = max(8, int(1.5 * dump_ms / max(kern_ms, 1e-3)) + 4, nslices + 4)
    if world > 1:                                              # the same number of steps (and pushes) on every rank
        mt = torch.tensor([M], device=dev)
        dist.all_reduce(mt, op=dist.ReduceOp.MAX)
        M = int(mt)

    def m_steps(with_dump):
        barrier()
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(stream):
            launch_step(d_in[0], 0, True)                      # untimed; its result seeds the dump
            a0.record(stream)
            if with_dump:
                dump_device()
            for j in range(M):                                  # j = 0 opens the next integration in the other span
                launch_step(d_in[j % NBUF], j % NBUF, j == 0, kept[1] if j == M - 1 else None)
            flush_dump()
            stream.wait_stream(gstream)
            if world > 1:
                stream.wait_stream(pstream)
            a1.record(stream)
        barrier()
        return a0.elapsed_time(a1)

    # the same M steps with no exchange at all (same launch, the fast dump stays on the GPU), before AND after the runs
    # with exchanges: a GPU under sw_power_cap slows by several per cent over the first ~100 ms of load, and a
    # one-sided comparison would charge that drift to the pushes
    def alone_steps():
        if not fused_fast:
            return None
        barrier()
        b0, b1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(stream):
            xc.accumulate_subsel(d_in[0], acc(), True, scratch_fast, True)
            b0.record(stream)
            for j in range(M):
                xc.accumulate_subsel(d_in[j % NBUF], acc(), False, scratch_fast, True)
            b1.record(stream)
        barrier()
        return b0.elapsed_time(b1)                           # (that integration is scrap: the next run opens a fresh one)

    scratch_fast = torch.zeros_like(fast[0])
    t_alone_before = alone_steps()
    t_plain = m_steps(False)
    t_dump = m_steps(True)
    dump_overlapped_ms = max(0.0, t_dump - t_plain)
    t_alone_after = alone_steps() if dumped[0] is not acc() else None
    t_alone = None if t_alone_before is None else 0.5 * (t_alone_before + (t_alone_after or t_alone_before))
    gather_ok = push_ok = None
    if world > 1:                                           # what arrived on rank 0 is what each rank sent
        with torch.cuda.stream(pstream):
            fpush.finish()
        torch.cuda.synchronize(dev)
        last = seqc[0] - 1                                   # the most recent step of all (the dump-overlap runs above)
        sent_dump = dumped[0]
        probe = torch.stack([sent_dump[:4096].to(torch.int64).sum(), sent_dump[-4096:].to(torch.int64).sum(),
                             kept[1].to(torch.int64).sum(), kept[1][-1].to(torch.int64).sum()])
        sent = [torch.empty_like(probe) for _ in range(world)]
        dist.all_gather(sent, probe)
        if rank == 0:
            got = peer.parts() if peer is not None else gather_list
            gather_ok = all(bool((torch.stack([got[r][:4096].to(torch.int64).sum(),
                                               got[r][-4096:].to(torch.int64).sum()]) == sent[r][:2]).all())
                            for r in range(world))
            if peer is not None and nslices > 0:             # ... and every rank's last slice is in (sequence words)
                gather_ok = gather_ok and bool((peer.arrived() == dseq[0] % 4096).all())
            payload, seqs = fpush.slot_on_root(last)
            payload = payload.view(world, nchan // NSUM, NVIS, 2)
            push_ok = bool((seqs == last).all()) and all(
                bool((torch.stack([payload[r].to(torch.int64).sum(), payload[r][-1].to(torch.int64).sum()])
                      == sent[r][2:]).all()) for r in range(world))
            if not gather_ok or not push_ok:
                raise RuntimeError("dump / fast dumps on rank 0 differ from what the ranks sent")

    # ---- e2e: host (pinned) voltages in, dump to pinned host memory out, copies inside the timed region
    h_in = [torch.empty((ntime, nchan, NINPUT), dtype=torch.uint8).pin_memory() for _ in range(2)]
    for k, h in enumerate(h_in):
        h.copy_(d_in[k].cpu())
    h_out = torch.empty(out_words, dtype=torch.int32).pin_memory()
    ring = [torch.empty_like(d_in[0]) for _ in range(3)]
    copy_stream = torch.cuda.Stream(device=dev)
    filled = [torch.cuda.Event() for _ in ring]
    freed = [torch.cuda.Event() for _ in ring]

    # the PCIe ceiling, on the record: plain pinned cudaMemcpyAsync H2D of the same gulp, nothing else running
    barrier()
    with torch.cuda.stream(copy_stream):
        p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ring[0].copy_(h_in[0], non_blocking=True)
        p0.record(copy_stream)
        for j in range(8):
            ring[j % 3].copy_(h_in[j % 2], non_blocking=True)
        p1.record(copy_stream)
    barrier()
    h2d_probe_gbs = 8 * in_bytes / p0.elapsed_time(p1) / 1e6

    h_view = h_out.view(2, nchan, wpc)
    h_nch = [nchan]                                             # channels of the last host dump

    def e2e_steps(n, first_index):
        d2h = 0
        for j in range(n):
            i = first_index + j
            slot = i % len(ring)
            with torch.cuda.stream(copy_stream):               # the Copy block: H2D into the gpu-input ring
                copy_stream.wait_event(freed[slot])
                ring[slot].copy_(h_in[i % 2], non_blocking=True)
                filled[slot].record(copy_stream)
            with torch.cuda.stream(stream):
                stream.wait_event(filled[slot])
                launch_step(ring[slot], i % 2, (i % accg) == 0)
                freed[slot].record(stream)
                due = (i % accg) == accg - 1
                if due or j == n - 1:
                    # CorrAcc's dump into a cuda_host ring span, side stream.  A dump that falls due leaves whole.  At the
                    # end of the region the running integration is read back in proportion to the gulps it holds
                    # (whole channels, rounded up): a full 383 MB read per 20-step run would put 12 x the bytes of
                    # the true cadence (one dump per acc_gulps steps) on a link that H2D and D2H share
                    nch_out = nchan if due else max(1, -(-nchan * ((i % accg) + 1) // accg))
                    src = acc().view(2, nchan, wpc)
                    stream.wait_stream(gstream)
                    gstream.wait_stream(stream)
                    with torch.cuda.stream(gstream):
                        for pl in range(2):
                            h_view[pl, :nch_out].copy_(src[pl, :nch_out], non_blocking=True)
                    h_nch[0] = nch_out
                    d2h += 2 * nch_out * wpc * 4
        return d2h

    for s_ in range(len(ring)):
        freed[s_].record(stream)
    n_e2e_warm = max(1, min(Wm, 6))
    e2e_steps(n_e2e_warm, 0)                                   # i = 0 opens a fresh integration
    barrier()
    t0 = torch.cuda.Event(enable_timing=True)
    t1 = torch.cuda.Event(enable_timing=True)
    t0.record(stream)
    copy_stream.wait_event(t0)
    d2h_total = e2e_steps(K, n_e2e_warm)
    stream.wait_stream(copy_stream)
    stream.wait_stream(gstream)
    if world > 1:
        stream.wait_stream(pstream)
    t1.record(stream)
    barrier()
    e2e_ms = t0.elapsed_time(t1)
    h_nch_timed = h_nch[0]
    clocks = sampler.stop() if sampler else None
    # the device->host result is really read, and it is the right one (same bytes through the H2D ring)
    h_acc = h_out.numpy().reshape(2, nchan, wpc)
    want = np.zeros((2, len(vchans), wpc), dtype=np.int64)
    for k in range(2):
        if counts[k]:
            v = np.ascontiguousarray(gulps_host[k][:, vchans]).reshape(ntime, len(vchans), NSTAND, NPOL)
            want += counts[k] * orc.corr_xgpu(v).astype(np.int64)
    want = want.astype(np.int32)
    in_dump = [n_ for n_, c_ in enumerate(vchans) if c_ < h_nch[0]]     # channel 0 is in every dump
    e2e_verified = bool(np.array_equal(h_acc[:, [vchans[n_] for n_ in in_dump]], want[:, in_dump]))
    # ... and, read back whole now (untimed), every checked channel of the accumulator the e2e region left behind
    h_out.copy_(acc())
    e2e_verified = e2e_verified and bool(np.array_equal(h_out.numpy().reshape(2, nchan, wpc)[:, vchans], want))
    if not e2e_verified:
        raise RuntimeError("e2e dump (pinned host) differs from the oracle on the same voltages")
    checksum = int(h_out[:1024].to(torch.int64).sum())

    # ---- whole X-engine (row N1) on the same process group
    full = None
    if not args.no_xengine_full:
        full = xengine_full(torch, dist, ops, sharding, orc, args, dev, world, rank, xc, d_in, gulps_host, acc(), fast,
                            vm, cj, stream, h_in, barrier)

    # ---- max over ranks (+ every rank's own per-step figures on rank 0)
    amort_ms = 0.0 if ndump > 0 else dump_overlapped_ms * K / accg   # no long dump fell in the timed region: charge its share
    per_rank = None
    if world > 1:
        t = torch.tensor([total_ms, e2e_ms, kern_ms, dump_ms, h2d_probe_gbs, dump_overlapped_ms, t_alone or 0.0,
                          t_plain, t_dump, t_alone_before or 0.0, t_alone_after or t_alone_before or 0.0],
                         dtype=torch.float64, device=dev)
        allt = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(allt, t)
        sv = torch.tensor(step_ms, dtype=torch.float64, device=dev)
        allsv = [torch.empty_like(sv) for _ in range(world)]
        dist.all_gather(allsv, sv)
        per_rank = [{"rank": r, "total_ms": float(allt[r][0]), "e2e_ms": float(allt[r][1]),
                     "step_ms": stats(allsv[r].cpu().numpy()),
                     "steps_ms": [round(float(x), 4) for x in allsv[r].cpu().numpy()],
                     "h2d_probe_gbs": float(allt[r][4]),
                     "e2e_h2d_gbs": in_bytes * K / float(allt[r][1]) / 1e6} for r in range(world)]
        m = torch.stack(allt).max(dim=0).values
        total_ms, e2e_ms, kern_ms, dump_ms = [float(x) for x in m[:4].tolist()]
        dump_overlapped_ms = float(m[5])
        t_alone, t_plain, t_dump, ta_b, ta_a = float(m[6]), float(m[7]), float(m[8]), float(m[9]), float(m[10])
        amort_ms = 0.0 if ndump > 0 else dump_overlapped_ms * K / accg

    if rank == 0:
        hbm_peak, peak_src = peaks()
        # int8 tensor peak measured in this run with a library GEMM (denominator only, never the product path)
        int8_meas = None
        try:
            a = torch.randint(-8, 8, (8192, 8192), dtype=torch.int8, device=dev)
            b = torch.randint(-8, 8, (8192, 8192), dtype=torch.int8, device=dev)
            for _ in range(3):
                torch._int_mm(a, b)
            torch.cuda.synchronize(dev)
            best = 1e9
            for _ in range(5):
                s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                s0.record()
                torch._int_mm(a, b)
                s1.record()
                torch.cuda.synchronize(dev)
                best = min(best, s0.elapsed_time(s1))
            int8_meas = 2 * 8192 ** 3 / best / 1e9
            del a, b
        except Exception:
            pass
        alg_bytes = in_bytes + 2 * out_words * 4                      # read gulp + read-modify-write accumulator
        if fused_fast:
            alg_bytes += 2 * fast[0].numel() * 4                      # fast dump zeroed + written
        traffic, traffic_src = ncu_traffic(nchan, ntime)
        achieved = alg_bytes / kern_ms / 1e6
        useful_tops = OPS_PER_SAMPLE * nsamp_step / kern_ms / 1e9
        eff_ms = total_ms + amort_ms
        value = world * nsamp_step * K / eff_ms / 1e6
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": Wm,
            "ms_per_step": eff_ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "int8->int32", "data": "synthetic",
            "config": workload_config(args),
            "clocks": clocks,
            "verified": {"accumulator_vs_oracle_same_bytes": verified, "channels": vchans,
                         "fast_dump_vs_oracle": fast_verified, "e2e_dump_vs_oracle": e2e_verified,
                         "how": "after the timed region the accumulator it left behind (and the pinned-host dump of "
                                "the e2e run) is compared word for word with oracle corr_xgpu + corr_acc on the same "
                                "voltages, on these channels"},
            "e2e": {"value": world * nsamp_step * K / e2e_ms / 1e6, "unit": UNIT,
                    "h2d_bytes_per_step": in_bytes, "d2h_bytes_per_step": d2h_total / K,
                    "ms_per_step": e2e_ms / K, "checksum": checksum,
                    "d2h_accounting": "a long dump that falls due inside the region goes to pinned host memory whole; at "
                                      "the end of the region the running integration is read back in proportion to the "
                                      "gulps it holds (%d of %d channels here), i.e. at the true cadence of one dump per "
                                      "acc_gulps steps; what was read is compared with the oracle, and so is the whole "
                                      "accumulator afterwards (untimed)" % (h_nch_timed, nchan),
                    "h2d_gbs_achieved_per_gpu": in_bytes * K / e2e_ms / 1e6,
                    "h2d_probe_gbs": h2d_probe_gbs,
                    "note": "h2d_probe_gbs = plain pinned cudaMemcpyAsync of the same gulp with nothing else running "
                            "(the PCIe ceiling of this box); the e2e rate is input bytes / time, 1 B per sample"},
            "gpu_launches": K * (1 + push_launches),
            "gpu_launches_whole_run": nlaunch[0],
            "step_ms": stats(step_ms),
            "dump": {"in_timed_region": ndump, "every_steps": accg, "alone_ms": dump_ms,
                     "overlapped_cost_ms": dump_overlapped_ms, "gather_verified": gather_ok,
                     "amortised_ms_added_per_step": amort_ms / K,
                     "gather": None if world == 1 else ("NCCL dist.gather" if args.nccl_gather else
                                                        "peer writes into rank 0's symmetric-memory buffer (copy engines over NVLink)"),
                     "bytes_to_rank0": (world - 1) * out_words * 4,
                     "slices": nslices,
                     "note": "dumps fall at their true cadence (every acc_gulps steps); alone_ms is one dump "
                             "(the finished span's copy to its consumer: N = 1 a device copy, N > 1 the gather to rank "
                             "0; side stream, the correlator accumulates the next integration in the other span of "
                             "the output ring) timed with nothing to overlap; overlapped_cost_ms is "
                             "what a dump adds to a run of steps it travels beside (side stream, copy engines; N > 1: "
                             "in `slices` pieces, one behind each of the next steps), "
                             "measured as (M steps + 1 dump) - (M steps), max over ranks; when no dump falls inside "
                             "the timed region its share overlapped_cost_ms * steps / acc_gulps is ADDED to the "
                             "timed total before value and ms_per_step are formed"},
            "roofline": {"kernel": "corr_swz_kernel (fused Corr+CorrAcc%s)" % ("+CorrSubsel" if fused_fast else ""),
                         "bound": "hbm", "achieved": achieved,
                         "peak": hbm_peak, "unit": "GB/s", "frac": achieved / hbm_peak, "traffic": traffic,
                         "traffic_source": traffic_src,
                         "peak_source": peak_src, "alg_bytes_per_launch": alg_bytes, "kernel_ms": kern_ms,
                         "kernel_ms_is": "median over the timed steps of the CUDA-event interval around one launch",
                         "reduce_path": {"payload_gbs": out_words * 4 / kern_ms / 1e6,
                                         "ceiling_gbs": 2600.0, "frac": out_words * 4 / kern_ms / 1e6 / 2600.0,
                                         "ceiling_source": "profiles/r03o_bulk_pattern.txt: cp.reduce.async.bulk.add.s32 "
                                                           "from 148 SMs into a > L2 footprint, wall clock; the fused "
                                                           "accumulate's read-modify-write goes through this path"},
                         "tensor": {"useful_int8_tops": useful_tops, "nominal_peak_tops": INT8_NOMINAL_TOPS,
                                    "frac_nominal": useful_tops / INT8_NOMINAL_TOPS,
                                    "measured_int8_gemm_tops": int8_meas,
                                    "frac_measured": (useful_tops / int8_meas) if int8_meas else None}},
            "realtime": {"required_gsps_per_%dch" % nchan: nchan * NINPUT * 23925.78125 / 1e9,
                         "full_band_required_gsps": 3072 * NINPUT * 23925.78125 / 1e9},
        }
        if world > 1:
            line["fast_dump_push"] = {"every_step": True, "bytes_per_rank_per_step": fast[0].numel() * 4,
                                      "how": push_kind, "verified_on_rank0": push_ok, "in_timed_region": K}
            line["exchange_overhead"] = {
                "steps": M, "ms_per_step_no_exchange": t_alone / M,
                "ms_per_step_no_exchange_before_after": [ta_b / M, ta_a / M], "ms_per_step_with_pushes": t_plain / M,
                "ms_per_step_with_pushes_and_one_dump": t_dump / M,
                "note": "the same M launches (fused Corr+CorrAcc+CorrSubsel, 384 ch) with the fast dump left on the GPU, "
                        "pushed to rank 0 every step, and pushed while one long dump travels; max over ranks.  The "
                        "no-exchange run is made before and after the other two: under sw_power_cap the same launch "
                        "slows by several per cent over the first ~100 ms of load, which is larger than anything the "
                        "exchanges add"}
            line["per_rank"] = per_rank
        if full is not None:
            line["xengine_full"] = full
        if not args.no_path_kernels and world == 1:
            del d_in, d_dump, ring
            torch.cuda.empty_cache()
            line["path_kernels"] = path_kernels(torch, ops, dev, nchan, hbm_peak)
        if not args.no_xengine_full and world == 1:
            try:
                line["blocks_graph"] = blocks_graph(torch, nchan)
            except Exception as e:                                      # a diagnostic section must not sink the line
                line["blocks_graph"] = {"error": repr(e)}
        if not args.no_cpu_baseline:
            v, desc, cores, _, _ = cpu_corr_sample(nchan * world, ntime, args.cpu_seconds, wrap=True, gulp=gulps_host[0])
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": desc}
            line["cpu_baseline_literal"] = cpu_literal_sample(gulps_host[0])
        real_stdout.write(json.dumps(line) + "\n")
        real_stdout.flush()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def xengine_full(torch, dist, ops, sharding, orc, args, dev, world, rank, xc, d_in, gulps_host, d_acc, fast, vm, cj,
                 stream, h_in, barrier):
    """Row N1: the WHOLE X-engine on every GPU -- per 480-sample gulp, fused Corr+CorrAcc+CorrSubsel on one stream and
    Beamform + power integration (32 beams, ntime_sum = 24, tensor-core kernel, weights pre-loaded) on the same bytes
    on a second stream, as production runs Corr and Beamform off the same gpu-input ring
    (lwa352-pipeline.py:232-234, 279-285).  N > 1: both per-step products (fast dump, beam powers) are pushed to rank 0.
    Timed device-resident and end to end (pinned host gulps through a 3-slot H2D ring shared by both consumers)."""
    nchan, ntime, K = args.nchan_per_gpu, args.ntime, args.full_steps
    nbeam, nsum = 32, 24
    nsamp = ntime * nchan * NINPUT
    bstream = torch.cuda.Stream(device=dev)
    bf = ops.Beamformer(gpu=dev.index, ninput=NINPUT, nchan=nchan, ntime=ntime, nbeam=nbeam)
    bf.set_stream(bstream)
    gw = torch.Generator(device=dev)
    gw.manual_seed(7 + rank)
    w = torch.randn((nchan, nbeam, NINPUT, 2), dtype=torch.float32, device=dev, generator=gw)
    with torch.cuda.stream(bstream):
        bf.set_weights(w)
    pw = [torch.zeros((nbeam // 2, ntime // nsum, nchan, 4), dtype=torch.float32, device=dev) for _ in range(2)]
    fpush = ppush = None
    pstream = torch.cuda.Stream(device=dev)
    if world > 1:
        try:
            fpush = sharding.PeerStream(fast[0].numel(), torch.int32, dev, dist, nslot=2, root=0, via=args.push)
            ppush = sharding.PeerStream(pw[0].numel(), torch.float32, dev, dist, nslot=2, root=0, via=args.push)
            ok = 1
        except Exception:
            ok = 0
        agreed = torch.tensor([ok], device=dev)
        dist.all_reduce(agreed, op=dist.ReduceOp.MIN)
        if int(agreed) == 0:
            fpush = sharding.GatherStream(fast[0].numel(), torch.int32, dev, dist, nslot=2, root=0)
            ppush = sharding.GatherStream(pw[0].numel(), torch.float32, dev, dist, nslot=2, root=0)
    a_done = [torch.cuda.Event() for _ in range(2)]
    b_done = [torch.cuda.Event() for _ in range(2)]
    slot_free = [torch.cuda.Event() for _ in range(2)]
    nlaunch = [0]

    def step(src, seq, first):
        s = seq % 2
        stream.wait_event(slot_free[s])
        bstream.wait_event(slot_free[s])
        with torch.cuda.stream(stream):
            xc.accumulate_subsel(src, d_acc, first, fast[s], True)
            a_done[s].record(stream)
        with torch.cuda.stream(bstream):
            bf.run_integrate(src, pw[s], None, nsum)
            b_done[s].record(bstream)
        nlaunch[0] += 2
        with torch.cuda.stream(pstream):
            pstream.wait_event(a_done[s])
            if fpush is not None:
                fpush.push(fast[s], seq)
                nlaunch[0] += 2 if getattr(fpush, "via", None) == "sm" else 0
            pstream.wait_event(b_done[s])
            if ppush is not None:
                ppush.push(pw[s], seq)
                nlaunch[0] += 2 if getattr(ppush, "via", None) == "sm" else 0
            slot_free[s].record(pstream)

    def run(n, srcs, e2e, reset):
        ring = filled = freed = copy_stream = None
        if e2e:
            ring = [torch.empty_like(d_in[0]) for _ in range(3)]
            copy_stream = torch.cuda.Stream(device=dev)
            filled = [torch.cuda.Event() for _ in ring]
            freed = [torch.cuda.Event() for _ in ring]
        barrier()
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0.record(stream)
        bstream.wait_event(t0)
        if e2e:
            copy_stream.wait_event(t0)
        for i in range(n):
            if e2e:
                r = i % 3
                with torch.cuda.stream(copy_stream):
                    copy_stream.wait_event(freed[r])
                    ring[r].copy_(srcs[i % len(srcs)], non_blocking=True)
                    filled[r].record(copy_stream)
                stream.wait_event(filled[r])
                bstream.wait_event(filled[r])
                step(ring[r], i, reset and i == 0)
                with torch.cuda.stream(pstream):            # both consumers are done with the ring slot
                    freed[r].record(pstream)
            else:
                step(srcs[i % len(srcs)], i, reset and i == 0)
        stream.wait_stream(bstream)
        stream.wait_stream(pstream)
        t1.record(stream)
        barrier()
        return t0.elapsed_time(t1)

    run(4, d_in, False, True)                                # warm-up (both kernels, the pushes); opens an integration
    dev_ms = run(K, d_in, False, False)
    # parity of what was just timed: the last step's fast dump (first output channel) and beam powers (two channels)
    last = K - 1
    v4 = np.ascontiguousarray(gulps_host[last % len(d_in)][:, :4]).reshape(ntime, 4, NSTAND, NPOL)
    ok_fast = bool(np.array_equal(fast[last % 2][0].cpu().numpy(), orc.subselect(orc.corr_xgpu(v4), vm, cj, 4)[0]))
    vb = np.ascontiguousarray(gulps_host[last % len(d_in)][:, :2])
    wb = w[:2].cpu().numpy()
    wantp = orc.beam_power_sum(orc.beamform(vb, wb[..., 0] + 1j * wb[..., 1]), nsum)
    gotp = pw[last % 2][:, :, :2].cpu().numpy()
    ok_pow = bool(np.abs(gotp - wantp).max() / np.abs(wantp).max() <= 1e-5)
    if not ok_fast or not ok_pow:
        raise RuntimeError("xengine_full: fast dump / beam powers differ from the oracle (%s, %s)" % (ok_fast, ok_pow))
    run(3, h_in, True, True)
    e2e_ms = run(K, h_in, True, False)
    pushed_ok = None
    if world > 1:
        with torch.cuda.stream(pstream):
            fpush.finish()
        torch.cuda.synchronize(dev)
        probe = torch.stack([fast[last % 2].to(torch.float64).sum(), pw[last % 2].to(torch.float64).sum()])
        sent = [torch.empty_like(probe) for _ in range(world)]
        dist.all_gather(sent, probe)
        if rank == 0:
            pf, sf = fpush.slot_on_root(last)
            pp, sp = ppush.slot_on_root(last)
            pushed_ok = bool((sf == last).all()) and bool((sp == last).all()) and all(
                float(pf[r].to(torch.float64).sum()) == float(sent[r][0]) and
                float(pp[r].to(torch.float64).sum()) == float(sent[r][1]) for r in range(world))
        t = torch.tensor([dev_ms, e2e_ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dev_ms, e2e_ms = float(t[0]), float(t[1])
    bf.close()
    full_band = 3072 * NINPUT * 23925.78125 / 1e9
    val = world * nsamp * K / dev_ms / 1e6
    e2e = world * nsamp * K / e2e_ms / 1e6
    return {
        "workload": "per step (one %d-sample gulp of %d channels x 704 inputs per GPU): fused Corr+CorrAcc+CorrSubsel "
                    "(4704 baselines, nchan_sum 4) and Beamform+power (32 beams, ntime_sum 24, weights pre-loaded) on "
                    "the same bytes, two streams%s" % (ntime, nchan, "; fast dump + beam powers pushed to rank 0 "
                                                       "every step" if world > 1 else ""),
        "steps": K, "ms_per_step": dev_ms / K, "gsamples_per_s": val, "per_gpu_gsamples_per_s": val / world,
        "e2e_ms_per_step": e2e_ms / K, "e2e_gsamples_per_s": e2e, "gpu_launches": nlaunch[0],
        "full_band_required_gsps": full_band, "x_realtime_full_band": val / full_band,
        "e2e_x_realtime_full_band": e2e / full_band,
        "verified": {"fast_dump_vs_oracle": ok_fast, "beam_powers_vs_oracle_1e-5": ok_pow, "pushes_on_rank0": pushed_ok},
        "push_bytes_per_rank_per_step": (fast[0].numel() * 4 + pw[0].numel() * 4) if world > 1 else 0,
    }


def blocks_graph(torch, nchan):
    """The same work driven through the Python blocks on the ring shim (tests/test_gpu_blocks.py's graph at full width):
    source -> Corr(acc_len = 480, slow_acc_len, fast dump) and Beamform(ntime_sum) on one input ring."""
    from xengine import blockbench
    return blockbench.run(nchan=nchan)


def main():
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
