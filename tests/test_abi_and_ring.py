"""CPU-side checks of the drop-in boundary: libxeng.so loads and exports every symbol include/xeng.h declares,
argument errors are reported through return codes (never exceptions / crashes), the library refuses to compute
without a GPU, and the ring shim behaves like the slice of the bifrost Ring API the blocks rely on."""
import ctypes as C
import json
import os
import re
import threading

import numpy as np
import pytest

from xengine import _lib, ops, ringshim

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    text = open(os.path.join(ROOT, "include", "xeng.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(xeng_[a-z0-9_]+)\s*\(", text)))


def test_every_declared_symbol_is_exported_and_bound():
    lib = _lib.load()
    names = header_symbols()
    assert len(names) >= 25
    for n in names:
        assert hasattr(lib, n), "libxeng.so does not export %s" % n
    assert set(names) == set(_lib.SIGNATURES), "ctypes SIGNATURES out of sync with include/xeng.h"
    assert lib.xeng_version() >= 100


def test_header_cites_the_reference_call_it_replaces():
    text = open(os.path.join(ROOT, "include", "xeng.h")).read()
    for needle in ("corr_block.py:253", "corr_block.py:445", "corr_block.py:331", "corr_acc_block.py:299",
                   "corr_subsel_block.py:298", "corr_output_full_block.py:669", "beamform_block.py:251",
                   "beamform_block.py:449", "beamform_sum_beams_block.py:245", "beamform_sum_single_beam_block.py:114"):
        assert needle in text, needle


def test_argument_errors_are_status_codes():
    lib = _lib.load()
    h = C.c_void_p()
    assert lib.xeng_corr_create(C.byref(h), 352, 3, 192, 480, -1) == -1          # npol != 2
    assert b"npol" in lib.xeng_last_error()
    assert lib.xeng_corr_create(C.byref(h), 350, 2, 192, 480, -1) == -1          # nstand not a multiple of 4
    assert lib.xeng_corr_create(None, 352, 2, 192, 480, -1) == -1
    assert lib.xeng_corr_run(None, None, None, 0) == -1
    assert lib.xeng_corr_matlen(None) == 0
    assert lib.xeng_acc(None, None, 16, 1, None) == -1
    assert lib.xeng_subsel(None, None, None, None, 1, 1, 1, 1, None) == -1
    assert lib.xeng_bf_run(None, None, None, None) == -1
    assert lib.xeng_bf_sync(None) == -1
    a2i = np.arange(8, dtype=np.int32)
    out = np.zeros(4 * 4 * 4, dtype=np.int32)
    assert lib.xeng_get_order(a2i.ctypes.data, out.ctypes.data, out.ctypes.data, 4, 3) == -1
    with pytest.raises(_lib.XengError):
        _lib.check(-1, "x")


def test_no_cpu_fallback_without_a_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    lib = _lib.load()
    h = C.c_void_p()
    rc = lib.xeng_corr_create(C.byref(h), 352, 2, 192, 480, -1)
    assert rc == -2 and b"no CUDA device" in lib.xeng_last_error() and not h.value
    rc = lib.xeng_bf_create(C.byref(h), -1, 704, 192, 480, 16, 0)
    assert rc == -2
    with pytest.raises(_lib.XengError):
        ops.XCorr(nstand=8, nchan=1, ntime_gulp=16)
    with pytest.raises(RuntimeError):
        ringshim.alloc_bytes(16, "cuda")


def test_missing_library_fails_loudly(monkeypatch):
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", "/nonexistent/libxeng.so")
    with pytest.raises(_lib.XengError, match="no CPU fallback"):
        _lib.load()


def test_get_order_is_a_permutation_onto_valid_words():
    nstand = 16
    a2i = np.arange(nstand * 2, dtype=np.int32).reshape(nstand, 2)[::-1].copy()   # a non-identity wiring
    a2bl, conj = ops.get_order(a2i)
    wpc = ops.words_per_chan(nstand)
    assert a2bl.min() >= 0 and a2bl.max() < wpc
    # (s0,p0)x(s1,p1) and (s1,p1)x(s0,p0) share a word with opposite conjugation, except for autos of one stand
    for s0 in range(nstand):
        for s1 in range(nstand):
            if s0 != s1:
                assert np.array_equal(a2bl[s0, s1], a2bl[s1, s0].T)
                assert np.array_equal(conj[s0, s1], 1 - conj[s1, s0].T)
            else:           # one stand: XX, XY, YX, YY are four distinct stored words, none conjugated on read
                assert len(np.unique(a2bl[s0, s0])) == 4 and conj[s0, s0].all()
    assert len(np.unique(a2bl[np.triu_indices(nstand)])) == nstand * (nstand + 1) // 2 * 4


# ---------------------------------------------------------------------------------------------- ring shim
def test_ring_rechunks_and_reports_short_final_gulp():
    ring = ringshim.Ring("r", space="system")
    with ring.begin_writing() as w:
        with w.begin_sequence(time_tag=5, header=json.dumps({"a": 1}), nringlet=1) as seq:
            for k in range(5):
                with seq.reserve(6) as span:
                    span.data[...] = np.arange(6 * k, 6 * k + 6, dtype=np.uint8)
    (rseq,) = list(ring.read())
    assert json.loads(rseq.header.tostring()) == {"a": 1} and rseq.time_tag == 5 and rseq.nringlet == 1
    spans = list(rseq.read(8))
    assert [s.size for s in spans] == [8, 8, 8, 6]
    assert np.array_equal(np.concatenate([np.asarray(s.data) for s in spans]), np.arange(30, dtype=np.uint8))
    assert spans[0].data_view('i32').shape == (2,)


def test_ring_two_readers_and_backpressure():
    ring = ringshim.Ring("r", space="system", maxspans=2)
    got = {0: [], 1: []}
    ready = threading.Barrier(3)

    def reader(k):
        it = ring.read(guarantee=True)
        ready.wait()
        for seq in it:
            for span in seq.read(4):
                got[k].append(int(np.asarray(span.data_view('i32'))[0]))

    threads = [threading.Thread(target=reader, args=(k,)) for k in (0, 1)]
    for t in threads:
        t.start()
    ready.wait()
    with ring.begin_writing() as w:
        for s in range(2):
            seq = w.begin_sequence(time_tag=s, header="{}")
            for k in range(10):
                span = ringshim.WriteSpan(seq.ring, 4)
                span.data_view('i32')[0] = 100 * s + k
                span.close()
            seq.end()
    for t in threads:
        t.join(timeout=20)
        assert not t.is_alive()
    want = [100 * s + k for s in range(2) for k in range(10)]
    assert got[0] == want and got[1] == want
