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


def _one(torch, nchan, ntime, ngulp, gulps_per_span, with_beam, slow_gulps, fast_space="cuda"):
    from . import ringshim
    from .blocks import Beamform, Corr
    log = logging.getLogger("blockbench")
    log.setLevel(logging.ERROR)
    dev = torch.device("cuda")
    span_gulps = max(gulps_per_span, 2 if with_beam else 1)          # Beamform reads 960-sample gulps
    gen = torch.Generator(device=dev)
    gen.manual_seed(1)
    bufs = [torch.randint(0, 256, (span_gulps * ntime * nchan * NINPUT,), dtype=torch.uint8, device=dev, generator=gen)
            for _ in range(3)]
    iring = ringshim.Ring("gpu-input", space="cuda", maxspans=4)
    slow = ringshim.Ring("corr-slow-output", space="cuda", recycle=True)
    fast = ringshim.Ring("corr-fast-output", space=fast_space, recycle=True)
    bfo = ringshim.Ring("bf-pow-output", space="cuda", recycle=True)
    corr = Corr(log, iring, None, ntime_gulp=ntime, nchan=nchan, nstand=NSTAND, npol=NPOL, acc_len=ntime, autostartat=0,
                ant_to_input=np.arange(NINPUT).reshape(NSTAND, NPOL), fast_oring=fast, acc_oring=slow,
                slow_acc_len=slow_gulps * ntime, gulps_per_span=gulps_per_span)
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

    def source():
        with iring.begin_writing() as w:
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
    return {"gulps": done, "ms_per_gulp": 1e3 * dt / done, "gsamples_per_s": done * ntime * nchan * NINPUT / dt / 1e9,
            "spans_out": {k: v[0] for k, v in counters.items()}}


def run(nchan=192, ntime=480, ngulp=240):
    import torch
    res = {"workload": "%d gulps of %d samples x %d channels x 704 inputs through Corr.main (fused CorrAcc + fast dump, "
                       "one launch per gulp)%s on the ring shim, device-resident input, wall clock over the whole run "
                       "(thread start to last span drained)" % (ngulp, ntime, nchan, " / + Beamform.main (32 beams, power mode)")}
    _one(torch, nchan, ntime, 24, 1, True, 12)                       # warm-up: kernels, allocator, threads
    res["corr_1_gulp_per_span"] = _one(torch, nchan, ntime, ngulp, 1, False, ngulp // 2)
    res["corr_6_gulps_per_span"] = _one(torch, nchan, ntime, ngulp, 6, False, ngulp // 2)
    res["corr_6_gulps_per_span_fast_ring_cuda_host"] = _one(torch, nchan, ntime, ngulp, 6, False, ngulp // 2, "cuda_host")
    res["corr_and_beamform"] = _one(torch, nchan, ntime, ngulp, 6, True, ngulp // 2)
    return res


if __name__ == "__main__":
    print(json.dumps(run(), indent=1))
