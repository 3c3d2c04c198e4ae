"""End-to-end parity of the block graph on the GPU, in the style of the reference's pipeline test scripts
(verification/tests/correlator/test_pipeline.sh -> lwa352-pipeline.py --fakesource; verification/tests/
beamformer/beamformer_test.py, beamformer_sum_test.py): a source thread writes 4+4-bit voltages into the
``gpu-input`` ring, every block runs ``main()`` on its own thread, the sinks compare with the numpy oracle.

    gpu-input -> Corr -> corr-output -+-> CorrAcc    -> corr-slow-output (cuda_host)
                                      +-> CorrSubsel -> corr-fast-output (cuda_host)
    gpu-input -> Beamform -> bf-output -> BeamformSumBeams -> bf-pow-output (cuda_host)

Integer products must be bit-exact; beam voltages / powers within 1e-5 relative (north_star).
"""
import json
import logging
import threading

import numpy as np
import pytest

from oracle import xeng_oracle as orc

pytestmark = pytest.mark.gpu
LOG = logging.getLogger("test")


@pytest.fixture(scope="module")
def torch_cuda():
    import torch
    if not torch.cuda.is_available():
        pytest.fail("gpu-marked test run without a CUDA device")
    return torch


def header(nchan, nstand, npol, seq0=0):
    return {'seq0': seq0, 'nchan': nchan, 'nstand': nstand, 'npol': npol, 'sfreq': 40.0e6, 'bw_hz': 23925.78125 * nchan,
            'chan0': 1672, 'fs_hz': 196000000, 'system_nchan': 3072, 'sync_time': 1, 'pipeline_id': 0,
            'ant_to_input': np.arange(nstand * npol).reshape(nstand, npol).tolist(),
            'input_to_ant': [[i // npol, i % npol] for i in range(nstand * npol)]}


def source(ring, hdr, gulps):
    """DummySource stand-in (dummy_source_block.py:221-262): one sequence, one span per gulp."""
    def run():
        with ring.begin_writing() as w:
            with w.begin_sequence(time_tag=0, header=json.dumps(hdr)) as seq:
                for g in gulps:
                    with seq.reserve(g.size) as span:
                        import torch
                        span.data.copy_(torch.from_numpy(np.ascontiguousarray(g).reshape(-1)))
    return threading.Thread(target=run, name="source")


def sink(ring, gulp, store):
    def run():
        for seq in ring.read(guarantee=True):
            hdr = json.loads(seq.header.tostring())
            spans = []
            for sp in seq.read(gulp):
                d = sp.data
                spans.append(d.cpu().numpy().copy() if hasattr(d, "cpu") else np.array(d))
            store.append((hdr, spans))
    return threading.Thread(target=run, name="sink-" + ring.name)


def run_all(threads, timeout=120):
    errors = []

    def guard(t):
        fn = t.run

        def wrapped():
            try:
                fn()
            except BaseException as e:          # surface exceptions of block threads in the test
                errors.append((t.name, repr(e)))
                raise
        t.run = wrapped
    for t in threads:
        guard(t)
        t.daemon = True
        t.start()
    for t in threads:
        t.join(timeout)
    assert not errors, errors
    assert not [t.name for t in threads if t.is_alive()], "pipeline did not drain"


def test_correlator_graph_matches_oracle(torch_cuda):
    from xengine import ringshim
    from xengine.blocks import Corr, CorrAcc, CorrSubsel
    nstand, npol, nchan, nt = 32, 2, 8, 32
    corr_acc_len, long_acc_len, ngulp = 2 * nt, 4 * nt, 13
    rng = np.random.default_rng(0xdeadbeef)
    v = rng.integers(0, 256, size=(ngulp, nt, nchan, nstand, npol), dtype=np.uint8)
    v[0].reshape(-1)[:4] = [0x88, 0x80, 0x08, 0xFF]
    r_in = ringshim.Ring("gpu-input", space="cuda")
    r_corr = ringshim.Ring("corr-output", space="cuda")
    r_slow = ringshim.Ring("corr-slow-output", space="cuda_host")
    r_fast = ringshim.Ring("corr-fast-output", space="cuda_host")
    hdr = header(nchan, nstand, npol)
    corr = Corr(LOG, r_in, r_corr, ntime_gulp=nt, nchan=nchan, npol=npol, nstand=nstand, acc_len=corr_acc_len,
                gpu=0, test=True, autostartat=nt, ant_to_input=np.array(hdr['ant_to_input']))
    acc = CorrAcc(LOG, r_corr, r_slow, nchan=nchan, npol=npol, nstand=nstand, acc_len=long_acc_len, gpu=0,
                  autostartat=nt + corr_acc_len)
    sub = CorrSubsel(LOG, r_corr, r_fast, nchan=nchan, npol=npol, nstand=nstand, nchan_sum=4, gpu=0,
                     antpol_to_bl=corr.antpol_to_bl, bl_is_conj=corr.bl_is_conj)
    bls = [[[int(a), int(b)], [int(c), int(d)]] for a, b, c, d in
           zip(rng.integers(0, nstand, sub.nvis_out), rng.integers(0, 2, sub.nvis_out),
               rng.integers(0, nstand, sub.nvis_out), rng.integers(0, 2, sub.nvis_out))]
    sub.process_command_strings(json.dumps({"id": "1", "cmd": "update", "val": {"kwargs": {"baselines": bls}}}))
    slow, fast = [], []
    run_all([source(r_in, hdr, list(v)), threading.Thread(target=corr.main, name="Corr"),
             threading.Thread(target=acc.main, name="CorrAcc"), threading.Thread(target=sub.main, name="CorrSubsel"),
             sink(r_slow, acc.ogulp_size, slow), sink(r_fast, sub.ogulp_size, fast)])

    # Corr integrations: gulps (1,2), (3,4), ... (11,12) -> 6 spans starting at sample nt
    want_corr = [orc.corr_xgpu(v[1 + 2 * k]) + orc.corr_xgpu(v[2 + 2 * k]) for k in range(6)]
    assert corr.stats.get('test_match') is True
    # fast path: every Corr integration, selected baselines, 4-channel sums
    (fh, fspans), = fast
    assert fh['nchan'] == nchan // 4 and fh['nvis'] == sub.nvis_out and fh['nchan_sum'] == 4 and fh['baselines'] == bls
    assert fh['seq0'] == nt and fh['acc_len'] == corr_acc_len
    vm, cj = orc.baselines_to_vismap(bls, corr.antpol_to_bl, corr.bl_is_conj)
    assert len(fspans) == 6
    for k, sp in enumerate(fspans):
        got = sp.view(np.int32).reshape(nchan // 4, sub.nvis_out, 2)
        assert np.array_equal(got, orc.subselect(want_corr[k].astype(np.int32), vm, cj, 4)), "fast dump %d" % k
    # the selected words really are v(s0,p0) conj(v(s1,p1)) summed over time and 4 channels (test_corr_part_rx.py:76-84)
    parts = [orc.corr_full(v[g]) for g in (1, 2)]
    full_re, full_im = parts[0][0] + parts[1][0], parts[0][1] + parts[1][1]
    got0 = fspans[0].view(np.int32).reshape(nchan // 4, sub.nvis_out, 2)
    for n in (0, 1, 17, sub.nvis_out - 1):
        (s0, p0), (s1, p1) = bls[n]
        assert got0[0, n, 0] == full_re[0:4, s0, s1, p0, p1].sum()
        assert got0[0, n, 1] == full_im[0:4, s0, s1, p0, p1].sum()
    # slow path: CorrAcc starts at the second Corr integration and sums 2 of them per dump
    (sh, sspans), = slow
    assert sh['upstream_acc_len'] == corr_acc_len and sh['acc_len'] == long_acc_len and sh['seq0'] == nt + corr_acc_len
    assert len(sspans) == 2
    for k, sp in enumerate(sspans):
        want = want_corr[1 + 2 * k] + want_corr[2 + 2 * k]
        assert np.array_equal(sp.view(np.int32).reshape(want.shape), want.astype(np.int32)), "slow dump %d" % k


def test_correlator_graph_fused_fast_dump(torch_cuda):
    """BASELINE configs[4] at block level: Corr(fast_oring=...) emits what Corr -> CorrSubsel emits (same header keys,
    same bytes) from the single fused launch, alongside the unchanged slow path through CorrAcc."""
    from xengine import ringshim
    from xengine.blocks import Corr, CorrAcc
    nstand, npol, nchan, nt, nvis = 32, 2, 8, 32, 4704
    corr_acc_len, long_acc_len, ngulp = 2 * nt, 4 * nt, 13
    rng = np.random.default_rng(0xfeed)
    v = rng.integers(0, 256, size=(ngulp, nt, nchan, nstand, npol), dtype=np.uint8)
    r_in = ringshim.Ring("gpu-input", space="cuda")
    r_corr = ringshim.Ring("corr-output", space="cuda")
    r_slow = ringshim.Ring("corr-slow-output", space="cuda_host")
    r_fast = ringshim.Ring("corr-fast-output", space="cuda")
    hdr = header(nchan, nstand, npol)
    corr = Corr(LOG, r_in, r_corr, ntime_gulp=nt, nchan=nchan, npol=npol, nstand=nstand, acc_len=corr_acc_len,
                gpu=0, autostartat=nt, ant_to_input=np.array(hdr['ant_to_input']), fast_oring=r_fast, nchan_sum=4,
                fast_nvis=nvis)
    acc = CorrAcc(LOG, r_corr, r_slow, nchan=nchan, npol=npol, nstand=nstand, acc_len=long_acc_len, gpu=0,
                  autostartat=nt + corr_acc_len)
    bls = [[[int(a), int(b)], [int(c), int(d)]] for a, b, c, d in
           zip(rng.integers(0, nstand, nvis), rng.integers(0, 2, nvis), rng.integers(0, nstand, nvis),
               rng.integers(0, 2, nvis))]
    corr.update_fast_subsel(bls)
    slow, fast = [], []
    run_all([source(r_in, hdr, list(v)), threading.Thread(target=corr.main, name="Corr"),
             threading.Thread(target=acc.main, name="CorrAcc"),
             sink(r_slow, acc.ogulp_size, slow), sink(r_fast, corr.fast_gulp_size, fast)])
    want_corr = [orc.corr_xgpu(v[1 + 2 * k]) + orc.corr_xgpu(v[2 + 2 * k]) for k in range(6)]
    (fh, fspans), = fast
    assert fh['nchan'] == nchan // 4 and fh['nvis'] == nvis and fh['nchan_sum'] == 4 and fh['baselines'] == bls
    assert fh['seq0'] == nt and fh['acc_len'] == corr_acc_len
    assert fh['sfreq'] == (hdr['sfreq'] + 3 * hdr['bw_hz'] / nchan) / 4
    vm, cj = orc.baselines_to_vismap(bls, corr.antpol_to_bl, corr.bl_is_conj)
    assert len(fspans) == 6
    for k, sp in enumerate(fspans):
        got = sp.view(np.int32).reshape(nchan // 4, nvis, 2)
        assert np.array_equal(got, orc.subselect(want_corr[k].astype(np.int32), vm, cj, 4)), "fused fast dump %d" % k
    (sh, sspans), = slow
    assert len(sspans) == 2
    for k, sp in enumerate(sspans):
        want = want_corr[1 + 2 * k] + want_corr[2 + 2 * k]
        assert np.array_equal(sp.view(np.int32).reshape(want.shape), want.astype(np.int32)), "slow dump %d" % k


@pytest.mark.parametrize("gulps_per_span,fast_space", [(1, "cuda_host"), (3, "cuda_host"), (2, "cuda")])
def test_correlator_graph_single_launch_long_accumulation_and_fast_dump(torch_cuda, gulps_per_span, fast_space):
    """The headline path at block level: Corr(oring=None, acc_oring, slow_acc_len, fast_oring) -- ONE launch per gulp
    does CorrAcc's long accumulation and CorrSubsel's fast dump (lwa352-pipeline.py:232-276: Corr -> CorrAcc +
    CorrSubsel, same header keys on both output rings).  Fast ring in production's 'cuda_host' space (the kernel
    gathers into device scratch), several gulps per input span, 4704 baselines, against the oracle."""
    from xengine import ringshim
    from xengine.blocks import Corr
    nstand, npol, nchan, nt, nvis = 32, 2, 8, 32, 4704
    fast_len, slow_len, ngulp = nt, 4 * nt, 14
    rng = np.random.default_rng(0xabcd + gulps_per_span)
    v = rng.integers(0, 256, size=(ngulp, nt, nchan, nstand, npol), dtype=np.uint8)
    r_in = ringshim.Ring("gpu-input", space="cuda")
    r_slow = ringshim.Ring("corr-slow-output", space="cuda_host")
    r_fast = ringshim.Ring("corr-fast-output", space=fast_space)
    hdr = header(nchan, nstand, npol)
    corr = Corr(LOG, r_in, None, ntime_gulp=nt, nchan=nchan, npol=npol, nstand=nstand, acc_len=fast_len, gpu=0,
                autostartat=nt, ant_to_input=np.array(hdr['ant_to_input']), fast_oring=r_fast, nchan_sum=4,
                fast_nvis=nvis, acc_oring=r_slow, slow_acc_len=slow_len, gulps_per_span=gulps_per_span)
    bls = [[[int(a), int(b)], [int(c), int(d)]] for a, b, c, d in
           zip(rng.integers(0, nstand, nvis), rng.integers(0, 2, nvis), rng.integers(0, nstand, nvis),
               rng.integers(0, 2, nvis))]
    corr.update_fast_subsel(bls)
    slow, fast = [], []
    run_all([source(r_in, hdr, list(v)), threading.Thread(target=corr.main, name="Corr"),
             sink(r_slow, corr.ogulp_size, slow), sink(r_fast, corr.fast_gulp_size, fast)])
    x = [orc.corr_xgpu(v[g]) for g in range(ngulp)]
    vm, cj = orc.baselines_to_vismap(bls, corr.antpol_to_bl, corr.bl_is_conj)
    (fh, fspans), = fast
    assert fh['nvis'] == nvis and fh['nchan'] == nchan // 4 and fh['acc_len'] == fast_len and fh['seq0'] == nt
    assert len(fspans) == ngulp - 1                       # every gulp from the start time on
    for k, sp in enumerate(fspans):
        got = sp.view(np.int32).reshape(nchan // 4, nvis, 2)
        assert np.array_equal(got, orc.subselect(x[1 + k], vm, cj, 4)), "fast dump %d" % k
    (sh, sspans), = slow
    assert sh['upstream_acc_len'] == fast_len and sh['acc_len'] == slow_len and sh['seq0'] == nt
    assert 'ant_to_input' not in sh
    assert len(sspans) == 3                               # gulps 1-4, 5-8, 9-12; 13 is an unfinished integration
    for k, sp in enumerate(sspans):
        want = sum(x[1 + 4 * k + j].astype(np.int64) for j in range(4))
        assert np.array_equal(sp.view(np.int32).reshape(want.shape), want.astype(np.int32)), "slow dump %d" % k


def test_ingest_source_copy_corr_chain(torch_cuda):
    """SURVEY 8 f2 as blocks: DummySource (pinned host ring, the reference's ramp fixture) -> Copy (H2D, one async copy
    per gulp, committed as each lands) -> gpu-input ring -> Corr; and the on-device source (device ring) -> Corr.
    Both dumps equal the oracle on the ramp (dummy_source_block.py:194-195: byte = stand % 8)."""
    from xengine import ringshim
    from xengine.blocks import Copy, Corr, DummySource
    nstand, npol, nchan, nt, ngulp = 32, 2, 8, 32, 6
    want = None
    for on_device in (False, True):
        r_host = ringshim.Ring("capture", space="cuda_host")
        r_in = ringshim.Ring("gpu-input", space="cuda")
        r_corr = ringshim.Ring("corr-output", space="cuda_host" if on_device else "cuda")
        src = DummySource(LOG, r_in if on_device else r_host, ntime_gulp=nt, nchan=nchan, nstand=nstand, npol=npol,
                          target_throughput=0, ngulp=ngulp)
        corr = Corr(LOG, r_in, r_corr, ntime_gulp=nt, nchan=nchan, npol=npol, nstand=nstand, acc_len=3 * nt, gpu=0,
                    autostartat=0, ant_to_input=src.ant_to_input)
        out = []
        threads = [threading.Thread(target=src.main, name="DummySource"), threading.Thread(target=corr.main, name="Corr"),
                   sink(r_corr, corr.ogulp_size, out)]
        if not on_device:
            threads.append(threading.Thread(
                target=Copy(LOG, r_host, r_in, ntime_gulp=nt, nbyte_per_time=nchan * nstand * npol, gpu=0,
                            buffer_multiplier=2).main, name="Copy"))
        run_all(threads)
        (hdr, spans), = out
        assert hdr['nchan'] == nchan and hdr['acc_len'] == 3 * nt and len(spans) == 2
        if want is None:
            want = 3 * orc.corr_xgpu(src.test_data).astype(np.int64)
        for sp in spans:
            assert np.array_equal(sp.view(np.int32).reshape(want.shape), want.astype(np.int32))


@pytest.mark.parametrize("use_cor_fmt", [True, False])
def test_corr_output_full_packets(torch_cuda, use_cor_fmt):
    """gpu-input -> Corr -> CorrAcc -> (cuda_host ring) -> CorrOutputFull: every packet's payload equals the oracle's
    v(s0) conj(v(s1)) for that baseline in the packet's own axis order, packets come s0-major, headers carry the
    stands (corr_output_full_block.py:440-548)."""
    import struct
    from xengine import ringshim
    from xengine.blocks import Corr, CorrAcc, CorrOutputFull
    nstand, npol, nchan, nt, ngulp = 36, 2, 40, 32, 7
    rng = np.random.default_rng(0xc0ffee)
    v = rng.integers(0, 256, size=(ngulp, nt, nchan, nstand, npol), dtype=np.uint8)
    r_in = ringshim.Ring("gpu-input", space="cuda")
    r_corr = ringshim.Ring("corr-output", space="cuda")
    r_slow = ringshim.Ring("corr-slow-output", space="cuda_host")
    hdr = header(nchan, nstand, npol)
    corr = Corr(LOG, r_in, r_corr, ntime_gulp=nt, nchan=nchan, npol=npol, nstand=nstand, acc_len=nt, gpu=0,
                autostartat=nt, ant_to_input=np.array(hdr['ant_to_input']))
    acc = CorrAcc(LOG, r_corr, r_slow, nchan=nchan, npol=npol, nstand=nstand, acc_len=2 * nt, gpu=0,
                  autostartat=2 * nt)
    out = CorrOutputFull(LOG, r_slow, nchan=nchan, npol=npol, nstand=nstand, antpol_to_bl=corr.antpol_to_bl,
                         bl_is_conj=corr.bl_is_conj, use_cor_fmt=use_cor_fmt, gpu=0)
    packets = []
    out.sink = lambda h, p: packets.append((h, p.copy()))
    run_all([source(r_in, hdr, list(v)), threading.Thread(target=corr.main, name="Corr"),
             threading.Thread(target=acc.main, name="CorrAcc"), threading.Thread(target=out.main, name="CorrOutputFull")])
    nbl = nstand * (nstand + 1) // 2
    assert len(packets) == 2 * nbl                     # gulps (2,3) and (4,5): two complete long integrations
    for k in range(2):
        fre, fim = orc.corr_full(v[2 + 2 * k])
        f2 = orc.corr_full(v[3 + 2 * k])
        fre, fim = fre + f2[0], fim + f2[1]            # [chan, s0, s1, p0, p1]
        n = k * nbl
        for s0 in range(nstand):
            for s1 in range(s0, nstand):
                h, p = packets[n]
                if use_cor_fmt:
                    assert len(h) == 32 and struct.unpack(">hh", h[28:32]) == (s0 + 1, s1 + 1)
                    got = p.reshape(nchan, npol, npol, 2)
                    assert np.array_equal(got[..., 0], fre[:, s0, s1]) and np.array_equal(got[..., 1], fim[:, s0, s1])
                else:
                    assert struct.unpack(">2I", h[-8:]) == (s0, s1)
                    assert struct.unpack(">QQ2d4I", h[:-8])[1] == (2 + 2 * k) * nt
                    got = p.reshape(npol, npol, nchan, 2)
                    assert np.array_equal(got[..., 0], np.moveaxis(fre[:, s0, s1], 0, -1))
                    assert np.array_equal(got[..., 1], np.moveaxis(fim[:, s0, s1], 0, -1))
                n += 1


@pytest.mark.parametrize("use_cor_fmt", [True, False])
def test_corr_output_full_checkfile(torch_cuda, tmp_path, use_cor_fmt):
    """CorrOutputFull(checkfile=...): the reference's in-pipeline check of the whole chain's conjugation convention
    (corr_output_full_block.py:550-603) against a file in make_golden_inputs.py's format (complex128
    [time][chan][stand][stand][pol][pol] = sum_t v(s0,p0) conj(v(s1,p1)) per checkfile_acc_len samples): every baseline
    good on the right file, and a file with one conjugated baseline is caught."""
    from xengine import ringshim
    from xengine.blocks import Corr, CorrAcc, CorrOutputFull
    nstand, npol, nchan, nt, ngulp = 12, 2, 8, 32, 5
    rng = np.random.default_rng(0xf11e)
    v = rng.integers(0, 256, size=(ngulp, nt, nchan, nstand, npol), dtype=np.uint8)
    golden = np.zeros((ngulp, nchan, nstand, nstand, npol, npol), dtype=complex)
    for g in range(ngulp):
        fre, fim = orc.corr_full(v[g])
        golden[g] = fre + 1j * fim
    good_file, bad_file = tmp_path / "corr_good.dat", tmp_path / "corr_bad.dat"
    golden.tofile(str(good_file))
    broken = golden.copy()
    broken[:, :, 3, 7] = np.conj(broken[:, :, 3, 7])
    broken.tofile(str(bad_file))
    for path, expect_bad in ((good_file, 0), (bad_file, None)):
        r_in = ringshim.Ring("gpu-input", space="cuda")
        r_corr = ringshim.Ring("corr-output", space="cuda")
        r_slow = ringshim.Ring("corr-slow-output", space="cuda_host")
        hdr = header(nchan, nstand, npol)
        corr = Corr(LOG, r_in, r_corr, ntime_gulp=nt, nchan=nchan, npol=npol, nstand=nstand, acc_len=nt, gpu=0,
                    autostartat=0, ant_to_input=np.array(hdr['ant_to_input']))
        acc = CorrAcc(LOG, r_corr, r_slow, nchan=nchan, npol=npol, nstand=nstand, acc_len=2 * nt, gpu=0, autostartat=0)
        out = CorrOutputFull(LOG, r_slow, nchan=nchan, npol=npol, nstand=nstand, antpol_to_bl=corr.antpol_to_bl,
                             bl_is_conj=corr.bl_is_conj, use_cor_fmt=use_cor_fmt, gpu=0, checkfile=str(path),
                             checkfile_acc_len=nt)
        run_all([source(r_in, hdr, list(v)), threading.Thread(target=corr.main, name="Corr"),
                 threading.Thread(target=acc.main, name="CorrAcc"), threading.Thread(target=out.main, name="CorrOutputFull")])
        nrows = nstand * (nstand + 1) // 2 * npol * npol * 2
        if expect_bad == 0:
            assert out.stats['checkfile_bad'] == 0 and out.stats['checkfile_good'] == nrows
        else:       # the imaginary rows of baseline (3, 7), all four pol products
            assert out.stats['checkfile_bad'] == npol * npol and out.stats['checkfile_good'] == nrows - npol * npol


@pytest.mark.parametrize("use_cor_fmt,device_layout", [(False, True), (True, True), (False, False), (True, False)])
def test_corr_output_part_packets(torch_cuda, use_cor_fmt, device_layout):
    """gpu-input -> Corr -> CorrSubsel -> (cuda_host ring) -> CorrOutputPart, with the payload order produced on the
    device (payload_layout) or by the block's host transpose: same packets either way, each equal to the oracle's
    channel-summed v(s0,p0) conj(v(s1,p1)) (corr_output_part_block.py:346-401; test_corr_part_rx.py:76-84)."""
    import struct
    from xengine import ringshim
    from xengine.blocks import Corr, CorrSubsel, CorrOutputPart
    nstand, npol, nchan, nt, ngulp, ncs = 16, 2, 8, 32, 4, 2
    rng = np.random.default_rng(0xbead)
    v = rng.integers(0, 256, size=(ngulp, nt, nchan, nstand, npol), dtype=np.uint8)
    r_in = ringshim.Ring("gpu-input", space="cuda")
    r_corr = ringshim.Ring("corr-output", space="cuda")
    r_fast = ringshim.Ring("corr-fast-output", space="cuda_host")
    hdr = header(nchan, nstand, npol)
    corr = Corr(LOG, r_in, r_corr, ntime_gulp=nt, nchan=nchan, npol=npol, nstand=nstand, acc_len=nt, gpu=0,
                autostartat=nt, ant_to_input=np.array(hdr['ant_to_input']))
    layout = ('cor' if use_cor_fmt else 'py') if device_layout else None
    sub = CorrSubsel(LOG, r_corr, r_fast, nchan=nchan, npol=npol, nstand=nstand, nchan_sum=ncs, gpu=0,
                     antpol_to_bl=corr.antpol_to_bl, bl_is_conj=corr.bl_is_conj, payload_layout=layout)
    # all dual-pol baselines of the first 48 stands' worth of slots: 4704 = 48*49/2 * 4, stands taken modulo nstand
    bls = []
    for i in range(48):
        for j in range(i, 48):
            for p0 in range(2):
                for p1 in range(2):
                    bls.append([[i % nstand, p0], [j % nstand, p1]])
    assert len(bls) == sub.nvis_out
    sub.process_command_strings(json.dumps({"id": "1", "cmd": "update", "val": {"kwargs": {"baselines": bls}}}))
    out = CorrOutputPart(LOG, r_fast, use_cor_fmt=use_cor_fmt, nvis_per_packet=16, nchan_sum=ncs)
    packets = []
    out.sink = lambda h, p: packets.append((h, p))
    run_all([source(r_in, hdr, list(v)), threading.Thread(target=corr.main, name="Corr"),
             threading.Thread(target=sub.main, name="CorrSubsel"), threading.Thread(target=out.main, name="CorrOutputPart")])
    nco = nchan // ncs
    per_dump = sub.nvis_out // 4 if use_cor_fmt else sub.nvis_out // 16
    assert len(packets) == 3 * per_dump                 # gulps 1, 2, 3
    for k in range(3):
        fre, fim = orc.corr_full(v[1 + k])
        fre = fre.reshape(nco, ncs, nstand, nstand, 2, 2).sum(1)
        fim = fim.reshape(nco, ncs, nstand, nstand, 2, 2).sum(1)
        for n in (0, 1, per_dump // 2, per_dump - 1):
            h, p = packets[k * per_dump + n]
            if use_cor_fmt:
                got = np.frombuffer(p, dtype=np.int32).reshape(nco, 4, 2)
                vis = bls[4 * n:4 * n + 4]
                assert len(h) == 32
            else:
                got = np.frombuffer(p, dtype='>i4').reshape(16, nco, 2).transpose(1, 0, 2)
                vis = bls[16 * n:16 * n + 16]
                assert np.frombuffer(h[struct.calcsize(">QQ2d4I"):], dtype='>i4').reshape(16, 2, 2).tolist() == vis
                assert struct.unpack(">QQ2d4I", h[:struct.calcsize(">QQ2d4I")])[1] == (1 + k) * nt
            for m, ((s0, p0), (s1, p1)) in enumerate(vis):
                assert np.array_equal(got[:, m, 0], fre[:, s0, s1, p0, p1])
                assert np.array_equal(got[:, m, 1], fim[:, s0, s1, p0, p1])


def test_beamformer_graph_matches_oracle(torch_cuda):
    from xengine import ringshim
    from xengine.blocks import Beamform, BeamformSumBeams
    nstand, npol, nchan, nt, nbeam, ntime_sum, ngulp = 40, 2, 16, 120, 8, 24, 3     # beamformer_test.py:18-26 shape
    ninput = nstand * npol
    rng = np.random.default_rng(0xaabbccdd)
    v = rng.integers(0, 255, size=(ngulp, nt, nchan, ninput), dtype=np.uint8)
    hdr = header(nchan, nstand, npol)
    freqs = hdr['sfreq'] + (hdr['bw_hz'] / nchan) * np.arange(nchan)
    cal = ((3 * rng.random((nchan, nbeam, ninput)) + 4) + 1j * (4 * rng.random((nchan, nbeam, ninput)) + 5)).astype(np.complex64)
    delays = 12 * rng.random((nbeam, ninput))
    amps = 7 * rng.random((nbeam, ninput)) + 10
    r_in = ringshim.Ring("gpu-input", space="cuda")
    r_bf = ringshim.Ring("bf-output", space="cuda")
    r_pow = ringshim.Ring("bf-pow-output", space="cuda_host")
    bf = Beamform(LOG, r_in, r_bf, nchan=nchan, nbeam=nbeam, ninput=ninput, ntime_gulp=nt, gpu=0)
    bfs = BeamformSumBeams(LOG, r_bf, r_pow, nchan=nchan, ntime_gulp=nt, ntime_sum=ntime_sum, gpu=0)
    bf.freqs = freqs            # commands sent before the first sequence use the block's current frequency axis
    for b in range(nbeam):      # as BeamformControl does: per-input calibration gains, then delays/amplitudes
        for i in range(ninput):
            data = np.stack([cal[:, b, i].real, cal[:, b, i].imag], axis=1).reshape(-1).tolist()
            bf.process_command_strings(json.dumps({"id": "c", "cmd": "update", "val": {"kwargs": {"coeffs": {
                "type": "calgains", "input_id": i, "beam_id": b, "data": data}}}}))
        bf.process_command_strings(json.dumps({"id": "b", "cmd": "update", "val": {"kwargs": {"coeffs": {
            "type": "beamcoeffs", "beam_id": b, "data": {"delays": delays[b].tolist(), "amps": amps[b].tolist()}}}}}))
    volts, pows = [], []
    # the voltage sink is a second reader of bf-output (the position of the reference's VLBI tap)
    run_all([source(r_in, hdr, list(v)), threading.Thread(target=bf.main, name="Beamform"),
             threading.Thread(target=bfs.main, name="BeamformSumBeams"),
             sink(r_bf, nt * nchan * nbeam * 8, volts), sink(r_pow, (nt // ntime_sum) * nchan * (nbeam // 2) * 16, pows)])
    w = np.stack([orc.beam_weights(freqs, delays[b], amps[b], cal[:, b, :]) for b in range(nbeam)], axis=1)
    assert np.allclose(bf.gains_cpu, w.astype(np.complex64), rtol=1e-6)
    (vh, vspans), = volts
    (ph, pspans), = pows
    assert (vh['nbeam'], vh['nstand'], vh['npol'], vh['nbit'], vh['complex']) == (nbeam, nbeam, 1, 32, True)
    assert (ph['nbeam'], ph['npol'], ph['acc_len']) == (nbeam // 2, 2, ntime_sum)
    assert len(vspans) == ngulp and len(pspans) == ngulp
    for g in range(ngulp):
        want_v = orc.beamform(v[g], bf.gains_cpu)
        got_v = vspans[g].view(np.complex64).reshape(nchan, nbeam, nt)
        assert np.abs(got_v - want_v).max() / np.abs(want_v).max() <= 1e-5
        want_p = orc.beam_power_sum(want_v, ntime_sum)
        got_p = pspans[g].view(np.float32).reshape(nbeam // 2, nt // ntime_sum, nchan, 4)
        assert np.abs(got_p - want_p).max() / np.abs(want_p).max() <= 1e-5


def test_beam_output_blocks_on_the_beamformer_graph(torch_cuda):
    """Beamform -> {BeamformVlbiOutput (device span through xeng_bf_vlbi_tap), BeamformSumBeams -> BeamformOutput}:
    the VLBI rows are the first two voltage beams of the oracle in [time][chan][beam] order
    (beamform_vlbi_output_block.py:260-262) and every power packet is the oracle's [chan][4] row of its (beam, time
    step) (beamform_output_block.py:353-359)."""
    from xengine import ringshim
    from xengine import runtime as rt
    from xengine.blocks import Beamform, BeamformOutput, BeamformSumBeams, BeamformVlbiOutput
    nstand, npol, nchan, nt, nbeam, ntime_sum, ngulp = 24, 2, 8, 96, 6, 24, 2
    ninput = nstand * npol
    rng = np.random.default_rng(77)
    v = rng.integers(0, 256, size=(ngulp, nt, nchan, ninput), dtype=np.uint8)
    hdr = header(nchan, nstand, npol)
    r_in = ringshim.Ring("gpu-input", space="cuda")
    r_bf = ringshim.Ring("bf-output", space="cuda")
    r_pow = ringshim.Ring("bf-pow-output", space="cuda_host")
    bf = Beamform(LOG, r_in, r_bf, nchan=nchan, nbeam=nbeam, ninput=ninput, ntime_gulp=nt, gpu=0)
    w = (rng.standard_normal((nchan, nbeam, ninput)) + 1j * rng.standard_normal((nchan, nbeam, ninput))).astype(np.complex64)
    bf.gains_cpu[...] = w
    rt.copy_array(bf.gains_gpu, bf.gains_cpu)          # what Beamform.main does when a load time passes
    bfs = BeamformSumBeams(LOG, r_bf, r_pow, nchan=nchan, ntime_gulp=nt, ntime_sum=ntime_sum, gpu=0)
    vlbi = BeamformVlbiOutput(LOG, r_bf, ntime_gulp=nt)
    pout = BeamformOutput(LOG, r_pow, ntime_gulp=nt // ntime_sum, pipeline_idx=2, nchan=nchan, nbeam=nbeam // 2)
    rows, pkts = [], []
    vlbi.sink = lambda seq, p: rows.append((seq, np.frombuffer(p, dtype=np.complex64).copy()))
    pout.sink = lambda b, h, p: pkts.append((b, h, np.frombuffer(p, dtype=np.float32).copy()))
    vlbi.process_command_strings(json.dumps({"id": "1", "cmd": "update", "val": {"kwargs": {"dest_ip": "10.1.1.1"}}}))
    pout.process_command_strings(json.dumps({"id": "1", "cmd": "update", "val": {"kwargs": {"dest_ip": ["10.1.1.2"]}}}))
    run_all([source(r_in, hdr, list(v)), threading.Thread(target=bf.main, name="Beamform"),
             threading.Thread(target=bfs.main, name="BeamformSumBeams"),
             threading.Thread(target=vlbi.main, name="BeamformVlbiOutput"),
             threading.Thread(target=pout.main, name="BeamformOutput")])
    assert len(rows) == ngulp * nt and len(pkts) == ngulp * (nbeam // 2) * (nt // ntime_sum)
    k = 0
    for g in range(ngulp):
        want_v = orc.beamform(v[g], bf.gains_cpu)                       # [chan][beam][time]
        for t in range(nt):
            seq, r = rows[g * nt + t]
            assert seq == g * nt + t
            ref = want_v[:, 0:2, t]
            assert np.abs(r.reshape(nchan, 2) - ref).max() / np.abs(want_v).max() <= 1e-5
        want_p = orc.beam_power_sum(want_v, ntime_sum)                  # [beam/2][T][chan][4]
        for b in range(nbeam // 2):
            for T in range(nt // ntime_sum):
                pb, _, p = pkts[k]
                k += 1
                assert pb == b
                assert np.abs(p.reshape(nchan, 4) - want_p[b, T]).max() / np.abs(want_p).max() <= 1e-5
