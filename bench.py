#!/usr/bin/env python
"""X-engine benchmark (contract: see the task statement / DESIGN.md "Measurement").

Workload (BASELINE.json configs[1]): Corr + CorrAcc on one B200 -- 704 inputs x 192 channels,
480-sample gulps, accumulated to 250-gulp (~5 s of sky) dumps.  One step = one 480-sample gulp
(64 880 640 4+4-bit complex samples per GPU) through the fused correlate+accumulate kernel; every
`--acc-gulps` steps the integration is dumped (device copy for `value`, D2H to pinned memory for `e2e`).
N > 1: every rank owns its own channel slab (weak scaling, no data-path collective); at each dump the
accumulated visibilities are gathered to rank 0 over NCCL.

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
    p.add_argument("--nchan-per-gpu", type=int, default=192)
    p.add_argument("--ntime", type=int, default=480)
    p.add_argument("--acc-gulps", type=int, default=250)
    p.add_argument("--nccl-gather", action="store_true",
                   help="gather dumps with dist.gather (NCCL kernels) instead of copy-engine peer writes")
    p.add_argument("--cpu-seconds", type=float, default=15.0, help="budget for the cpu_baseline sample")
    p.add_argument("--no-cpu-baseline", action="store_true")
    p.add_argument("--no-path-kernels", action="store_true", help="skip the per-kernel figures of the rest of the path")
    return p.parse_args()


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
def cpu_corr_sample(nchan_total, ntime, budget_s, seed=0xdeadbeef, wrap=False):
    """Time oracle Corr (xGPU order) + CorrAcc on a bounded sample of the workload: a slab of channels of
    one gulp, grown until ~budget_s of CPU work has been spent (wrap=True carries on into further gulps when
    the host finishes one gulp inside the budget).  Returns (Gsamples/s, description, cores)."""
    from oracle import xeng_oracle as orc
    rng = np.random.RandomState(seed & 0xFFFFFFFF)
    nch = 2
    v = rng.randint(0, 255, size=(ntime, nch, NSTAND, NPOL), dtype=np.uint8)   # make_golden_inputs.py:137
    orc.corr_xgpu(v)                                                           # warm-up (BLAS threads, index maps)
    done_ch, t_used = 0, 0.0
    acc = None
    while t_used < budget_s and (wrap or done_ch < nchan_total):
        t0 = time.perf_counter()
        x = orc.corr_xgpu(v)
        acc = orc.corr_acc(x.copy() if acc is None else acc, x, first=acc is None)
        t_used += time.perf_counter() - t0
        done_ch += nch
    samples = done_ch * ntime * NINPUT
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    desc = "%d channels = %.2f gulps of %d channels x %d samples, 704 inputs (numpy/BLAS float32-exact matmul " \
           "port of Corr._test + CorrAcc, all host threads), %.1f s" % (done_ch, done_ch / nchan_total, nchan_total,
                                                                        ntime, t_used)
    return samples / t_used / 1e9, desc, cores, t_used, samples


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
        "workload": "Corr+CorrAcc: 704 inputs (352 stands x 2 pol) x %d channels/GPU, %d-sample gulps of 4+4-bit "
                    "voltages, fused correlate+accumulate, dump every %d gulps" % (args.nchan_per_gpu, args.ntime,
                                                                                 args.acc_gulps),
        "nchan_per_gpu": args.nchan_per_gpu, "ninput": NINPUT, "ntime_gulp": args.ntime,
        "acc_gulps": args.acc_gulps,
        "sharding": "channel slabs, 1 rank per GPU over torch.distributed/NCCL, each dump gathered to rank 0 on a side stream (overlaps the next integration, inside the timed region)" if args.gpus > 1 else "single GPU",
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
    for nvis in (4704, 4656):
        sel = np.random.default_rng(nvis).integers(0, a2bl.size, nvis)
        vm = torch.from_numpy(a2bl.reshape(-1)[sel].copy()).to(dev)
        cj = torch.from_numpy(conj.reshape(-1)[sel].copy()).to(dev)
        so = torch.zeros((nchan // 4, nvis, 2), dtype=torch.int32, device=dev)
        ms = timeit(lambda: ops.subsel(vis, so, vm, cj, nvis, nchan, 4, wpc))
        ab = 8 * nvis * nchan + 8 * nvis * nchan // 4
        res["corr_subsel_%d" % nvis] = {"ms": ms, "alg_bytes": ab, "gbs": ab / ms / 1e6,
                                        "sector_bytes": 2 * 32 * nvis * nchan + 8 * nvis * nchan // 4}
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
    sel = np.random.default_rng(1).integers(0, a2bl.size, 4704)
    xc.set_subsel(a2bl.reshape(-1)[sel].copy(), conj.reshape(-1)[sel].copy(), 4)
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
    del gulps, acc2, fast
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
    return res


# ----------------------------------------------------------------------------------------------
# B200 arm
# ----------------------------------------------------------------------------------------------
def run_b200(args):
    import torch
    import torch.distributed as dist
    from xengine import ops, sharding

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

    # synthetic voltages: 4 distinct gulps, uniform bytes, fixed seed (per rank = per channel slab)
    g = torch.Generator(device=dev)
    g.manual_seed(0xdeadbeef + rank)
    NBUF = 4
    d_in = [torch.randint(0, 256, (ntime, nchan, NINPUT), dtype=torch.uint8, device=dev, generator=g)
            for _ in range(NBUF)]
    d_acc = torch.zeros(out_words, dtype=torch.int32, device=dev)
    d_dump = torch.zeros(out_words, dtype=torch.int32, device=dev)
    gather_list = None
    stream = torch.cuda.Stream(device=dev)
    xc = ops.XCorr(nstand=NSTAND, npol=NPOL, nchan=nchan, ntime_gulp=ntime, gpu=local)
    xc.set_stream(stream)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    gstream = torch.cuda.Stream(device=dev) if world > 1 else None
    peer = None
    if world > 1 and not args.nccl_gather:
        try:
            peer = sharding.PeerGather(out_words, torch.int32, dev, dist, root=0)
        except Exception as e:                              # no symmetric memory on this node: all ranks use NCCL
            sys.stderr.write("rank %d: PeerGather unavailable (%s)\n" % (rank, e))
        agreed = torch.tensor([1 if peer is not None else 0], device=dev)
        dist.all_reduce(agreed, op=dist.ReduceOp.MIN)
        if int(agreed) == 0:
            peer, args.nccl_gather = None, True
        else:
            d_dump = peer.slot if rank == 0 else d_dump    # rank 0 snapshots straight into its own slot

    if world > 1 and rank == 0 and args.nccl_gather:
        gather_list = [d_dump] + [torch.empty_like(d_dump) for _ in range(world - 1)]

    def dump_device():
        # CorrAcc's copy_array into its output span, then (N > 1) the NCCL gather of that span to rank 0 on a side
        # stream: as in the pipeline, the correlator carries on with the next integration while the previous dump
        # travels; the next dump waits for it, and so does the end of the timed region.
        if world > 1:
            stream.wait_stream(gstream)
        d_dump.copy_(d_acc, non_blocking=True)
        if world > 1:
            gstream.wait_stream(stream)
            with torch.cuda.stream(gstream):
                if peer is not None:
                    peer.push(d_dump)
                else:
                    sharding.gather_to_root(d_dump, dist, root=0, out=gather_list)

    def step_device(i):
        xc.accumulate(d_in[i % NBUF], d_acc, (i % accg) == 0)
        if (i % accg) == accg - 1:
            dump_device()

    # ---- value: inputs resident in HBM
    with torch.cuda.stream(stream):
        for i in range(Wm):
            step_device(i)
        dump_device()          # warm the dump path too (first NCCL gather sets up its peer connections)
    barrier()
    sampler = ClockSampler(local) if rank == 0 else None
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(K + 1)]
    ndump = 0
    phase = accg // 2
    with torch.cuda.stream(stream):
        # the timed region starts in the middle of a long integration (steady state): a dump falls every `accg`
        # steps, the first after accg - accg // 2, and travels to rank 0 while the next integration runs
        evs[0].record(stream)
        for i in range(K):
            g_i = i + phase
            xc.accumulate(d_in[i % NBUF], d_acc, (g_i % accg) == 0)
            evs[i + 1].record(stream)
            if (g_i % accg) == accg - 1:
                dump_device()
                ndump += 1
        if world > 1:
            stream.wait_stream(gstream)
        end = torch.cuda.Event(enable_timing=True)
        end.record(stream)
    barrier()
    total_ms = evs[0].elapsed_time(end)
    # kernel-only time: intervals that contain exactly one corr kernel (those after a dump also contain the copy)
    kern = []
    for i in range(K):
        prev_dump = i > 0 and (((i - 1 + phase) % accg) == accg - 1)
        if not prev_dump:
            kern.append(evs[i].elapsed_time(evs[i + 1]))
    kern_ms = float(np.mean(kern)) if kern else total_ms / K
    xc.sync()
    # one dump on its own (copy + gather, nothing to overlap with): what each long integration adds at most
    barrier()
    with torch.cuda.stream(stream):
        dmp0 = torch.cuda.Event(enable_timing=True)
        dmp1 = torch.cuda.Event(enable_timing=True)
        dmp0.record(stream)
        dump_device()
        if world > 1:
            stream.wait_stream(gstream)
        dmp1.record(stream)
    barrier()
    dump_ms = dmp0.elapsed_time(dmp1)
    gather_ok = None
    if world > 1:                                           # what arrived on rank 0 is what each rank sent
        probe = torch.stack([d_dump[:4096].to(torch.int64).sum(), d_dump[-4096:].to(torch.int64).sum()])
        sent = [torch.empty_like(probe) for _ in range(world)]
        dist.all_gather(sent, probe)
        if rank == 0:
            got = peer.parts() if peer is not None else gather_list
            gather_ok = all(bool((torch.stack([got[r][:4096].to(torch.int64).sum(),
                                               got[r][-4096:].to(torch.int64).sum()]) == sent[r]).all())
                            for r in range(world))
            if not gather_ok:
                raise RuntimeError("gathered dump on rank 0 differs from what the ranks sent")

    # ---- e2e: host (pinned) voltages in, dump to pinned host memory out, copies inside the timed region
    h_in = [torch.empty((ntime, nchan, NINPUT), dtype=torch.uint8).pin_memory() for _ in range(2)]
    for k, h in enumerate(h_in):
        h.copy_(d_in[k].cpu())
    h_out = torch.empty(out_words, dtype=torch.int32).pin_memory()
    ring = [torch.empty_like(d_in[0]) for _ in range(3)]
    copy_stream = torch.cuda.Stream(device=dev)
    filled = [torch.cuda.Event() for _ in ring]
    freed = [torch.cuda.Event() for _ in ring]

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
                xc.accumulate(ring[slot], d_acc, (i % accg) == 0)
                freed[slot].record(stream)
                if (i % accg) == accg - 1 or j == n - 1:
                    h_out.copy_(d_acc, non_blocking=True)      # CorrAcc dump into a cuda_host ring span
                    d2h += out_words * 4
        return d2h

    for s in range(len(ring)):
        freed[s].record(stream)
    e2e_steps(min(Wm, 6), 0)
    barrier()
    t0 = torch.cuda.Event(enable_timing=True)
    t1 = torch.cuda.Event(enable_timing=True)
    t0.record(stream)
    copy_stream.wait_event(t0)
    d2h_total = e2e_steps(K, 0)
    stream.wait_stream(copy_stream)
    t1.record(stream)
    barrier()
    e2e_ms = t0.elapsed_time(t1)
    clocks = sampler.stop() if sampler else None
    checksum = int(h_out[:1024].to(torch.int64).sum())        # the device->host result is really read

    # ---- max over ranks
    if world > 1:
        t = torch.tensor([total_ms, e2e_ms, kern_ms, dump_ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms, e2e_ms, kern_ms, dump_ms = [float(x) for x in t.tolist()]

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
        traffic, traffic_src = ncu_traffic(nchan, ntime)
        achieved = alg_bytes / kern_ms / 1e6
        useful_tops = OPS_PER_SAMPLE * nsamp_step / kern_ms / 1e9
        value = world * nsamp_step * K / total_ms / 1e6
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": Wm,
            "ms_per_step": total_ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "int8->int32", "data": "synthetic",
            "config": workload_config(args),
            "clocks": clocks,
            "e2e": {"value": world * nsamp_step * K / e2e_ms / 1e6, "unit": UNIT,
                    "h2d_bytes_per_step": in_bytes, "d2h_bytes_per_step": d2h_total / K,
                    "ms_per_step": e2e_ms / K, "checksum": checksum},
            "gpu_launches": K,
            "dump": {"in_timed_region": ndump, "every_steps": accg, "alone_ms": dump_ms, "gather_verified": gather_ok,
                     "gather": None if world == 1 else ("NCCL dist.gather" if args.nccl_gather else
                                                        "peer writes into rank 0's symmetric-memory buffer (copy engines over NVLink)"),
                     "bytes_to_rank0": (world - 1) * out_words * 4,
                     "note": "dumps fall at their true cadence (every acc_gulps steps); alone_ms is one dump "
                             "(device copy + gather to rank 0) timed with nothing to overlap"},
            "roofline": {"kernel": "corr_swz_kernel (fused Corr+CorrAcc)", "bound": "hbm", "achieved": achieved,
                         "peak": hbm_peak, "unit": "GB/s", "frac": achieved / hbm_peak, "traffic": traffic,
                         "traffic_source": traffic_src,
                         "peak_source": peak_src, "alg_bytes_per_launch": alg_bytes, "kernel_ms": kern_ms,
                         "tensor": {"useful_int8_tops": useful_tops, "nominal_peak_tops": INT8_NOMINAL_TOPS,
                                    "frac_nominal": useful_tops / INT8_NOMINAL_TOPS,
                                    "measured_int8_gemm_tops": int8_meas,
                                    "frac_measured": (useful_tops / int8_meas) if int8_meas else None}},
            "realtime": {"required_gsps_per_%dch" % nchan: nchan * NINPUT * 23925.78125 / 1e9,
                         "full_band_required_gsps": 3072 * NINPUT * 23925.78125 / 1e9},
        }
        if not args.no_path_kernels and world == 1:
            del d_in, d_dump, ring
            torch.cuda.empty_cache()
            line["path_kernels"] = path_kernels(torch, ops, dev, nchan, hbm_peak)
        if not args.no_cpu_baseline:
            v, desc, cores, _, _ = cpu_corr_sample(nchan * world, ntime, args.cpu_seconds, wrap=True)
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": desc}
        real_stdout.write(json.dumps(line) + "\n")
        real_stdout.flush()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
