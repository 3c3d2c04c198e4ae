"""Throughput of the X-engine driven through the Python BLOCKS on the ring shim (bench.py's ``blocks_graph``).

The graph is production's (lwa352-pipeline.py:232-285) with the fused options on: one device input ring feeds
``Corr(acc_len = ntime_gulp, acc_oring = slow ring, slow_acc_len, fast_oring = fast ring)`` -- one launch per gulp does
the long accumulation and the fast dump -- and ``Beamform(ntime_sum = 24)``; three sinks drain the output rings.  Every
block runs ``main()`` on its own thread, as in the reference.  The source commits pre-generated device gulps (zero-copy),
so the figure isolates what the block layer costs on top of the raw C-ABI rate: Python bookkeeping per gulp, span
reservation, the waits for the device before spans are committed.
"""
import json
import logging
import threading
import time

import numpy as np

NSTAND, NPOL = 352, 2
NINPUT = NSTAND * NPOL


def _header(nchan, seq0=0):
    return {'seq0': seq0, 'nchan': nchan, 'nstand': NSTAND, 'npol': NPOL, 'sfreq': 30e6, 'bw_hz': 23925.78125 * nchan,
            'chan0': 0, 'fs_hz': 196000000, 'system_nchan': nchan, 'sync_time': 0, 'pipeline_id': 0}


def _drain(ring, gulp, counter):
    for seq in ring.read(guarantee=True):
        for span in seq.read(gulp):
            counter[0] += 1


def _one(torch, nchan, ntime, ngulp, gulps_per_span, with_beam, slow_gulps, fast_space="cuda", profile=False,
         from_host=False):
    from . import ringshim
    from .blocks import Beamform, Copy, Corr
    log = logging.getLogger("blockbench")
    log.setLevel(logging.ERROR)
    dev = torch.device("cuda")
    span_gulps = max(gulps_per_span, 2 if with_beam else 1)          # Beamform reads 960-sample gulps
    gen = torch.Generator(device=dev)
    gen.manual_seed(1)
    bufs = [torch.randint(0, 256, (span_gulps * ntime * nchan * NINPUT,), dtype=torch.uint8, device=dev, generator=gen)
            for _ in range(3)]
    iring = ringshim.Ring("gpu-input", space="cuda", maxspans=4, recycle=from_host)
    hring = None
    if from_host:           # capture ring in pinned host memory -> Copy block (H2D) -> gpu-input, as production does
        hring = ringshim.Ring("capture", space="cuda_host", maxspans=4)
        bufs = [b.cpu().pin_memory() for b in bufs]
    slow = ringshim.Ring("corr-slow-output", space="cuda", recycle=True)
    fast = ringshim.Ring("corr-fast-output", space=fast_space, recycle=True)
    bfo = ringshim.Ring("bf-pow-output", space="cuda", recycle=True)
    corr = Corr(log, iring, None, ntime_gulp=ntime, nchan=nchan, nstand=NSTAND, npol=NPOL, acc_len=ntime, autostartat=0,
                ant_to_input=np.arange(NINPUT).reshape(NSTAND, NPOL), fast_oring=fast, acc_oring=slow,
                slow_acc_len=slow_gulps * ntime,
                gulps_per_span=1 if from_host else gulps_per_span)     # the Copy block commits gulp-sized spans
    prof = None
    if profile:
        import cProfile
        prof = cProfile.Profile()
        threads = [threading.Thread(target=lambda: prof.runcall(corr.main))]
    else:
        threads = [threading.Thread(target=corr.main)]
    counters = {"slow": [0], "fast": [0], "bf": [0]}
    threads.append(threading.Thread(target=_drain, args=(slow, corr.ogulp_size, counters["slow"])))
    threads.append(threading.Thread(target=_drain, args=(fast, corr.fast_gulp_size, counters["fast"])))
    if with_beam:
        nbeam, nsum = 32, 24
        bf = Beamform(log, iring, bfo, nchan=nchan, nbeam=nbeam, ninput=NINPUT, ntime_gulp=2 * ntime, ntime_sum=nsum)
        bf.gains_cpu[...] = (np.random.default_rng(0).standard_normal(bf.gains_cpu.shape) + 1j).astype(np.complex64)
        threads.append(threading.Thread(target=bf.main))
        threads.append(threading.Thread(target=_drain, args=(bfo, (2 * ntime // nsum) * nchan * nbeam * 8, counters["bf"])))
    nspan = ngulp // span_gulps
    if from_host:
        cp = Copy(log, hring, iring, ntime_gulp=ntime, nbyte_per_time=nchan * NINPUT, buffer_multiplier=span_gulps)
        threads.append(threading.Thread(target=cp.main))

    def source():
        with (hring if from_host else iring).begin_writing() as w:
            with w.begin_sequence(time_tag=0, header=json.dumps(_header(nchan))) as seq:
                for k in range(nspan):
                    seq.commit_buffer(bufs[k % len(bufs)])

    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for t in threads:
        t.start()
    src = threading.Thread(target=source)
    src.start()
    src.join()
    for t in threads:
        t.join()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    done = nspan * span_gulps
    if prof is not None:
        import io
        import pstats
        buf = io.StringIO()
        pstats.Stats(prof, stream=buf).sort_stats("tottime").print_stats(14)
        import sys
        sys.stderr.write("---- profile gulps_per_span=%d fast=%s\n%s\n" % (gulps_per_span, fast_space, buf.getvalue()))
    return {"gulps": done, "ms_per_gulp": 1e3 * dt / done, "gsamples_per_s": done * ntime * nchan * NINPUT / dt / 1e9,
            "spans_out": {k: v[0] for k, v in counters.items()}}


def run(nchan=192, ntime=480, ngulp=240):
    import torch
    res = {"workload": "%d gulps of %d samples x %d channels x 704 inputs through Corr.main (fused CorrAcc + fast dump, "
                       "one launch per gulp)%s on the ring shim, device-resident input, wall clock over the whole run "
                       "(thread start to last span drained)" % (ngulp, ntime, nchan, " / + Beamform.main (32 beams, power mode)")}
    import gc

    def case(*a, **kw):
        best = None
        for _ in range(2):                                           # first pass warms kernels, allocator, threads
            gc.collect()
            r = _one(torch, nchan, ntime, *a, **kw)
            if best is None or r["ms_per_gulp"] < best["ms_per_gulp"]:
                best = r
        return best

    res["corr_1_gulp_per_span"] = case(ngulp, 1, False, ngulp // 2)
    res["corr_6_gulps_per_span"] = case(ngulp, 6, False, ngulp // 2)
    res["corr_6_gulps_per_span_fast_ring_cuda_host"] = case(ngulp, 6, False, ngulp // 2, "cuda_host")
    res["corr_and_beamform"] = case(ngulp, 6, True, ngulp // 2)
    res["e2e_pinned_host_via_copy_block_corr_and_beamform"] = case(ngulp // 2, 2, True, ngulp // 4, from_host=True)
    res["note"] = "best of two passes per case"
    return res


if __name__ == "__main__":
    import sys
    if len(sys.argv) > 1 and sys.argv[1] == "profile":
        import torch
        _one(torch, 192, 480, 24, 1, False, 12)
        for gps, space in ((1, "cuda"), (6, "cuda"), (6, "cuda_host")):
            print(gps, space, _one(torch, 192, 480, 120, gps, False, 60, space, profile=True))
    else:
        print(json.dumps(run(), indent=1))
