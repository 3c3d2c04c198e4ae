"""INTEGRATION.md option B, executed: the reference's UNMODIFIED block classes running on libxeng.

Two halves, because the reference sources exist only in the build container and the GPU only on the GPU box:

* ``test_reference_blocks_run_unmodified_over_the_adapter`` (CPU, needs /root/reference): imports the reference's own
  ``corr_block.py``, ``corr_acc_block.py``, ``corr_subsel_block.py``, ``beamform_block.py``,
  ``beamform_sum_beams_block.py`` with ``bifrost`` resolved to ``xengine/bfcompat`` (``_bf`` = ``XengBF``), runs their
  ``main()`` over ring-shim rings and compares every output span with the oracle.  Without a GPU the four libxeng
  entry points behind the adapter are replaced by oracle-backed doubles that work on the same raw addresses, so what
  this half proves is the call surface: argument order, array descriptors, dump / clear semantics, ring and header
  plumbing of the unmodified reference code over the adapter.
* ``test_adapter_call_sequence_on_the_gpu`` (gpu): the same ``_bf.*`` calls, in the order and with the arguments the
  reference blocks make them (file:line cited per call), on compat arrays in device memory with the REAL libxeng
  underneath, against the oracle.
"""
import ctypes as C
import importlib
import json
import logging
import os
import sys

import numpy as np
import pytest

from oracle import xeng_oracle as orc

REF = "/root/reference/pipeline"
HERE = os.path.dirname(os.path.abspath(__file__))
COMPAT = os.path.join(os.path.dirname(HERE), "caltech-bifrost-dsp_b200", "xengine", "bfcompat")
LOG = logging.getLogger("dropin")
NSTAND, NPOL, NCHAN, NT = 8, 2, 4, 16
NINPUT = NSTAND * NPOL


def _arr(addr, n, dtype):
    return np.ctypeslib.as_array((C.c_uint8 * (n * np.dtype(dtype).itemsize)).from_address(addr)).view(dtype)


class OracleXCorr(object):
    """ops.XCorr double: same calls, raw addresses in and out, arithmetic by the oracle."""

    def __init__(self, nstand=352, npol=2, nchan=192, ntime_gulp=480, gpu=-1, stream=None):
        self.n = (nstand, npol, nchan, ntime_gulp)
        self.wpc = orc.words_per_chan(nstand)
        self.acc = None

    def run(self, a_in, a_out, do_dump):
        ns, npol, nchan, nt = self.n
        v = _arr(a_in, nt * nchan * ns * npol, np.uint8).reshape(nt, nchan, ns, npol)
        x = orc.corr_xgpu(v).astype(np.int64)
        self.acc = x if self.acc is None else self.acc + x
        if do_dump:
            _arr(a_out, 2 * nchan * self.wpc, np.int32)[:] = self.acc.astype(np.int32).reshape(-1)
            self.acc = None

    def sync(self):
        pass


def oracle_subsel(a_in, a_out, a_vm, a_cj, nvis, nchan, nsum, wpc, stream=None):
    x = _arr(a_in, 2 * nchan * wpc, np.int32).reshape(2, nchan, wpc)
    vm, cj = _arr(a_vm, nvis, np.int32), _arr(a_cj, nvis, np.int32)
    _arr(a_out, (nchan // nsum) * nvis * 2, np.int32)[:] = orc.subselect(x, vm, cj, nsum).reshape(-1)


def oracle_acc(a, b, n, first, stream=None):
    a, b = np.asarray(a).reshape(-1), np.asarray(b).reshape(-1)
    if first:
        a[:n] = b[:n]
    else:
        a[:n] += b[:n]


class OracleBeamformer(object):
    def __init__(self, gpu=-1, ninput=704, nchan=192, ntime=480, nbeam=16, ntime_blocks=0, stream=None):
        self.ninput, self.nchan, self.ntime, self.nbeam = ninput, nchan, ntime, nbeam

    def run(self, a_in, a_out, a_w):
        v = _arr(a_in, self.ntime * self.nchan * self.ninput, np.uint8).reshape(self.ntime, self.nchan, self.ninput)
        w = _arr(a_w, self.nchan * self.nbeam * self.ninput, np.complex64).reshape(self.nchan, self.nbeam, self.ninput)
        _arr(a_out, self.nchan * self.nbeam * self.ntime, np.complex64)[:] = orc.beamform(v, w).astype(np.complex64).reshape(-1)

    def integrate(self, a_in, a_out, ntime_sum):
        x = _arr(a_in, self.nchan * self.nbeam * self.ntime, np.complex64).reshape(self.nchan, self.nbeam, self.ntime)
        p = orc.beam_power_sum(x, ntime_sum).astype(np.float32)
        _arr(a_out, p.size, np.float32)[:] = p.reshape(-1)

    def sync(self):
        pass


def _header(seq0=0):
    return {'seq0': seq0, 'nchan': NCHAN, 'nstand': NSTAND, 'npol': NPOL, 'sfreq': 1000.0, 'bw_hz': 100.0 * NCHAN,
            'chan0': 0, 'fs_hz': 196000000, 'system_nchan': 32 * NCHAN, 'sync_time': 0, 'pipeline_id': 0,
            'ant_to_input': np.arange(NINPUT).reshape(NSTAND, NPOL).tolist(),
            'input_to_ant': [[i // NPOL, i % NPOL] for i in range(NINPUT)]}


@pytest.fixture
def reference_blocks(host_only, monkeypatch):
    if not os.path.isdir(REF):
        pytest.skip("the reference sources live only in the build container")
    from xengine import ops
    monkeypatch.setattr(ops, "XCorr", OracleXCorr)
    monkeypatch.setattr(ops, "subsel", oracle_subsel)
    monkeypatch.setattr(ops, "acc", oracle_acc)
    monkeypatch.setattr(ops, "Beamformer", OracleBeamformer)
    if not hasattr(np, "complex"):
        monkeypatch.setattr(np, "complex", complex, raising=False)      # the reference predates numpy 1.24
    saved = {k: sys.modules.get(k) for k in list(sys.modules) if k.split(".")[0] in ("bifrost", "lwa352_pipeline", "ujson")}
    for k in saved:
        del sys.modules[k]
    monkeypatch.syspath_prepend(REF)
    monkeypatch.syspath_prepend(COMPAT)
    monkeypatch.setitem(sys.modules, "ujson", json)
    import bifrost.libbifrost as lb
    lb.configure(nstand=NSTAND, npol=NPOL, nchan=NCHAN, ntime_gulp=NT)
    mods = {n: importlib.import_module("lwa352_pipeline.blocks." + n)
            for n in ("corr_block", "corr_acc_block", "corr_subsel_block", "beamform_block", "beamform_sum_beams_block")}
    for m in mods.values():
        assert m.__file__.startswith(REF), m.__file__            # the reference's own files, untouched
    import bifrost.ring as bfring
    yield mods, bfring
    for k in [k for k in sys.modules if k.split(".")[0] in ("bifrost", "lwa352_pipeline")]:
        del sys.modules[k]
    sys.modules.update({k: v for k, v in saved.items() if v is not None})


def _fill(ring, hdr, gulps):
    with ring.begin_writing() as w:
        with w.begin_sequence(time_tag=0, header=json.dumps(hdr)) as seq:
            for g in gulps:
                with seq.reserve(g.size) as span:
                    np.asarray(span._span.data).reshape(-1)[:] = g.reshape(-1)


def _drain(ring, gulp, dtype):
    out = []
    for seq in ring._shim.read(guarantee=False):
        hdr = json.loads(seq.header.tostring())
        out.append((hdr, [np.array(np.asarray(sp.data).reshape(-1)[:sp.size]).view(dtype) for sp in seq.read(gulp)]))
    return out


def test_reference_blocks_run_unmodified_over_the_adapter(reference_blocks):
    mods, bfring = reference_blocks
    rng = np.random.default_rng(0xb1f)
    ngulp = 9
    v = rng.integers(0, 256, size=(ngulp, NT, NCHAN, NSTAND, NPOL), dtype=np.uint8)
    # ---- Corr -> CorrAcc + CorrSubsel (lwa352-pipeline.py:232-276)
    r_in, r_corr = bfring.Ring(space="cuda", name="gpu-input"), bfring.Ring(space="cuda", name="corr-output")
    r_slow, r_fast = bfring.Ring(space="cuda_host", name="slow"), bfring.Ring(space="cuda_host", name="fast")
    _fill(r_in, _header(), list(v))
    Corr, CorrAcc, CorrSubsel = mods["corr_block"].Corr, mods["corr_acc_block"].CorrAcc, mods["corr_subsel_block"].CorrSubsel
    corr = Corr(LOG, r_in, r_corr, ntime_gulp=NT, nchan=NCHAN, npol=NPOL, nstand=NSTAND, acc_len=2 * NT, gpu=-1,
                autostartat=NT, ant_to_input=np.arange(NINPUT).reshape(NSTAND, NPOL))
    a2bl, conj = orc.get_order(np.arange(NINPUT, dtype=np.int32).reshape(NSTAND, NPOL))
    assert np.array_equal(np.asarray(corr.antpol_to_bl), a2bl) and np.array_equal(np.asarray(corr.bl_is_conj), conj)
    corr.main()
    x = [orc.corr_xgpu(g).astype(np.int64) for g in v]
    want_corr = [x[1 + 2 * k] + x[2 + 2 * k] for k in range(4)]
    acc = CorrAcc(LOG, r_corr, r_slow, nchan=NCHAN, npol=NPOL, nstand=NSTAND, acc_len=4 * NT, gpu=-1, autostartat=NT)
    acc.main()
    sub = CorrSubsel(LOG, r_corr, r_fast, nchan=NCHAN, npol=NPOL, nstand=NSTAND, nchan_sum=2, gpu=-1,
                     antpol_to_bl=corr.antpol_to_bl, bl_is_conj=corr.bl_is_conj)
    bls = [[[int(a), int(b)], [int(c), int(d)]] for a, b, c, d in
           zip(rng.integers(0, NSTAND, sub.nvis_out), rng.integers(0, 2, sub.nvis_out),
               rng.integers(0, NSTAND, sub.nvis_out), rng.integers(0, 2, sub.nvis_out))]
    sub.process_command_strings(json.dumps({"id": "1", "cmd": "update", "val": {"kwargs": {"baselines": bls}}}))
    sub.main()
    (hdr, spans), = _drain(r_corr, corr.ogulp_size, np.int32)
    assert hdr['acc_len'] == 2 * NT and hdr['seq0'] == NT and len(spans) == 4
    for k, sp in enumerate(spans):
        assert np.array_equal(sp, want_corr[k].astype(np.int32).reshape(-1)), "Corr dump %d" % k
    (shdr, sspans), = _drain(r_slow, acc.ogulp_size, np.int32)
    assert shdr['upstream_acc_len'] == 2 * NT and shdr['acc_len'] == 4 * NT and len(sspans) == 2
    for k, sp in enumerate(sspans):
        assert np.array_equal(sp, (want_corr[2 * k] + want_corr[2 * k + 1]).astype(np.int32).reshape(-1))
    (fhdr, fspans), = _drain(r_fast, sub.ogulp_size, np.int32)
    vm, cj = orc.baselines_to_vismap(bls, a2bl, conj)
    assert fhdr['nvis'] == sub.nvis_out and fhdr['nchan_sum'] == 2 and len(fspans) == 4
    for k, sp in enumerate(fspans):
        assert np.array_equal(sp.reshape(NCHAN // 2, sub.nvis_out, 2), orc.subselect(want_corr[k].astype(np.int32), vm, cj, 2))
    # ---- Beamform -> BeamformSumBeams (lwa352-pipeline.py:279-285)
    nbeam, nsum = 4, 4
    Beamform, BfSum = mods["beamform_block"].Beamform, mods["beamform_sum_beams_block"].BeamformSumBeams
    r_in2, r_bf, r_pow = bfring.Ring(space="cuda"), bfring.Ring(space="cuda"), bfring.Ring(space="cuda_host")
    _fill(r_in2, _header(), [g.reshape(NT, NCHAN, NINPUT) for g in v[:3]])
    bf = Beamform(LOG, r_in2, r_bf, nchan=NCHAN, nbeam=nbeam, ninput=NINPUT, ntime_gulp=NT, gpu=-1)
    w = (rng.standard_normal((NCHAN, nbeam, NINPUT)) + 1j * rng.standard_normal((NCHAN, nbeam, NINPUT))).astype(np.complex64)
    bf.gains_cpu[...] = w
    bf.main()
    bs = BfSum(LOG, r_bf, r_pow, nchan=NCHAN, ntime_gulp=NT, ntime_sum=nsum, gpu=-1)
    bs.main()
    (bh, bspans), = _drain(r_bf, NT * NCHAN * nbeam * 8, np.complex64)
    assert bh['nbeam'] == nbeam and bh['nbit'] == 32 and len(bspans) == 3
    (ph, pspans), = _drain(r_pow, (NT // nsum) * NCHAN * (nbeam // 2) * 16, np.float32)
    assert ph['nbeam'] == nbeam // 2 and ph['acc_len'] == nsum and len(pspans) == 3
    for k in range(3):
        wantv = orc.beamform(v[k].reshape(NT, NCHAN, NINPUT), w)
        assert np.allclose(bspans[k].reshape(wantv.shape), wantv, rtol=1e-5, atol=1e-3)
        wantp = orc.beam_power_sum(wantv.astype(np.complex64), nsum)
        assert np.allclose(pspans[k].reshape(wantp.shape), wantp, rtol=1e-4, atol=1e-2)


def test_our_blocks_run_on_a_foreign_ring_type(host_only, monkeypatch):
    """INTEGRATION.md option A: THIS repo's blocks on rings that are not the shim -- the ``bifrost.ring.Ring`` of
    xengine/bfcompat, which exposes only bifrost's public ring API (no shim internals: spans come from
    ``bifrost.ring.WriteSpan``, span data are foreign array objects).  Host logic only (libxeng doubles)."""
    for k in [k for k in sys.modules if k.split(".")[0] == "bifrost"]:
        del sys.modules[k]
    monkeypatch.syspath_prepend(COMPAT)
    import bifrost.ring as bfring
    from xengine.blocks import Corr, CorrAcc
    r_in, r_corr, r_slow = bfring.Ring(space="cuda"), bfring.Ring(space="cuda"), bfring.Ring(space="cuda_host")
    for r in (r_in, r_corr, r_slow):
        assert not hasattr(r, "_commit") and not hasattr(r, "_open_seq")
    gulps = []
    for t in range(1, 8):
        g = np.zeros(NT * NCHAN * NSTAND * NPOL, dtype=np.uint8)
        g[0] = t
        gulps.append(g)
    _fill(r_in, _header(), gulps)
    corr = Corr(LOG, r_in, r_corr, ntime_gulp=NT, nchan=NCHAN, npol=NPOL, nstand=NSTAND, acc_len=2 * NT, autostartat=0,
                ant_to_input=np.arange(NINPUT).reshape(NSTAND, NPOL))
    corr.main()
    CorrAcc(LOG, r_corr, r_slow, nchan=NCHAN, npol=NPOL, nstand=NSTAND, acc_len=4 * NT, autostartat=0).main()
    (hdr, spans), = _drain(r_corr, corr.ogulp_size, np.int32)
    assert hdr['acc_len'] == 2 * NT and [tuple(s[:2]) for s in spans] == [(1 + 2, 2), (3 + 4, 2), (5 + 6, 2)]
    (shdr, sspans), = _drain(r_slow, corr.ogulp_size, np.int32)
    assert shdr['upstream_acc_len'] == 2 * NT and [int(s[0]) for s in sspans] == [1 + 2 + 3 + 4]
    for k in [k for k in sys.modules if k.split(".")[0] == "bifrost"]:
        del sys.modules[k]


@pytest.mark.gpu
def test_adapter_call_sequence_on_the_gpu():
    """The reference blocks' ``_bf.*`` call sequence on the real library (device memory, compat arrays)."""
    import torch
    if not torch.cuda.is_available():
        pytest.fail("gpu-marked test run without a CUDA device")
    sys.path.insert(0, COMPAT)
    try:
        import bifrost.ndarray as BFArray
        from bifrost.ndarray import copy_array
        from bifrost.device import stream_synchronize
        from xengine.bfcompat.adapter import XengBF
        nstand, npol, nchan, nt = 32, 2, 8, 64
        ninput = nstand * npol
        _bf = XengBF(nstand=nstand, npol=npol, nchan=nchan, ntime_gulp=nt)
        rng = np.random.default_rng(5)
        v = rng.integers(0, 256, size=(3, nt, nchan, nstand, npol), dtype=np.uint8)
        # corr_block.py:251-253
        assert _bf.bfXgpuInitialize(BFArray([0], dtype='i8', space='cuda').as_BFarray(),
                                    BFArray([0], dtype='i64', space='cuda').as_BFarray(), 0) == _bf.BF_STATUS_SUCCESS
        # corr_block.py:259-261, 329-333
        a2i = BFArray(np.zeros([nstand, npol], dtype=np.int32), space='system')
        a2bl = BFArray(np.zeros([nstand, nstand, npol, npol], dtype=np.int32), space='system')
        conj = BFArray(np.zeros([nstand, nstand, npol, npol], dtype=np.int32), space='system')
        a2i[...] = np.arange(ninput).reshape(nstand, npol)
        assert _bf.bfXgpuGetOrder(a2i.as_BFarray(), a2bl.as_BFarray(), conj.as_BFarray()) == 0
        wa, wc = orc.get_order(np.asarray(a2i))
        assert np.array_equal(np.asarray(a2bl), wa) and np.array_equal(np.asarray(conj), wc)
        # corr_block.py:445: one call per gulp, doDump on the last gulp of the integration; then a second integration
        matlen = nchan * orc.words_per_chan(nstand)
        ospan = BFArray(shape=[2 * matlen], dtype='i32', space='cuda')
        for integ in range(2):
            gulps = [0, 1] if integ == 0 else [2]
            for i, g in enumerate(gulps):
                ispan = BFArray(v[g].reshape(-1), space='cuda')
                assert _bf.bfXgpuKernel(ispan.as_BFarray(), ospan.as_BFarray(), int(i == len(gulps) - 1)) == 0
            want = sum(orc.corr_xgpu(v[g]).astype(np.int64) for g in gulps).astype(np.int32)
            assert np.array_equal(ospan.copy(space='system').reshape(want.shape), want), "integration %d" % integ
        # corr_subsel_block.py:210-215, 281-282, 298-305
        nvis = 4704
        bls = [[[int(a), int(b)], [int(c), int(d)]] for a, b, c, d in
               zip(rng.integers(0, nstand, nvis), rng.integers(0, 2, nvis), rng.integers(0, nstand, nvis), rng.integers(0, 2, nvis))]
        vm, cj = orc.baselines_to_vismap(bls, wa, wc)
        subsel, subsel_next = BFArray(shape=[nvis], dtype='i32', space='cuda'), BFArray(shape=[nvis], dtype='i32', space='cuda_host')
        cjd, cj_next = BFArray(shape=[nvis], dtype='i32', space='cuda'), BFArray(shape=[nvis], dtype='i32', space='cuda_host')
        for k in range(nvis):
            subsel_next.data[k] = int(vm[k])
            cj_next.data[k] = int(cj[k])
        copy_array(subsel, subsel_next)
        copy_array(cjd, cj_next)
        obuf = BFArray(shape=[nchan // 4, nvis], dtype='ci32', space='cuda')
        idata = ospan.view('ci32').reshape(matlen)
        assert _bf.bfXgpuSubSelect(idata.as_BFarray(), obuf.as_BFarray(), subsel.as_BFarray(), cjd.as_BFarray(), 4, 0) == 0
        odata = BFArray(shape=[nchan // 4, nvis], dtype='ci32', space='cuda_host')
        copy_array(odata, obuf)
        stream_synchronize()
        assert np.array_equal(np.asarray(odata.copy(space='system')), orc.subselect(want, vm, cj, 4))
        # beamform_block.py:241, 251-253, 449; beamform_sum_beams_block.py:220, 245
        nbeam, ntb, nsum = 8, 96, 24
        assert _bf.bfBeamformInitialize(0, ninput, nchan, ntb, nbeam, 0) == 0
        vb = rng.integers(0, 256, size=(ntb, nchan, ninput), dtype=np.uint8)
        w = (rng.standard_normal((nchan, nbeam, ninput)) + 1j * rng.standard_normal((nchan, nbeam, ninput))).astype(np.complex64)
        gains = BFArray(shape=(nchan, nbeam, ninput), dtype=np.complex64, space='cuda')
        copy_array(gains, w)
        ib = BFArray(vb.reshape(-1), space='cuda')
        ob = BFArray(shape=[nchan * nbeam * ntb * 2], dtype=np.float32, space='cuda')
        assert _bf.bfBeamformRun(ib.as_BFarray(), ob.as_BFarray(), gains.as_BFarray()) == 0
        bout = BFArray(shape=(nbeam // 2, ntb // nsum, nchan, 4), dtype=np.float32, space='cuda')
        assert _bf.bfBeamformIntegrate(ob.as_BFarray(), bout.as_BFarray(), nsum) == 0
        stream_synchronize()
        _bf.sync()
        wantv = orc.beamform(vb, w)
        gotv = np.asarray(ob.copy(space='system')).reshape(nchan, nbeam, ntb, 2)
        assert np.abs(gotv[..., 0] + 1j * gotv[..., 1] - wantv).max() / np.abs(wantv).max() <= 1e-5
        wantp = orc.beam_power_sum(wantv, nsum)
        assert np.abs(np.asarray(bout.copy(space='system')) - wantp).max() / np.abs(wantp).max() <= 1e-5
        # a failing call returns a status instead of raising (corr_block.py:254-256)
        assert _bf.bfBeamformIntegrateSingleBeam(ob.as_BFarray(), bout.as_BFarray(), nsum, 99) != 0
    finally:
        sys.path.remove(COMPAT)
        for k in [k for k in sys.modules if k.split(".")[0] == "bifrost"]:
            del sys.modules[k]
