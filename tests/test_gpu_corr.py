"""GPU parity of Corr / CorrAcc / CorrSubsel / reorder against the numpy oracle, through the C-ABI.
Bit-exact (integer work).  Mirrors pipeline/verification/xgpu_test.py and Corr._compare of the reference."""
import numpy as np
import pytest

from oracle import xeng_oracle as orc

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def torch_cuda():
    import torch
    if not torch.cuda.is_available():
        pytest.fail("gpu-marked test run without a CUDA device")
    return torch


def _voltages(rng, nt, nchan, nstand, npol=2, edge=True):
    v = rng.integers(0, 256, size=(nt, nchan, nstand, npol), dtype=np.uint8)
    if edge and v.size >= 8:  # bytes randint(0,255) never yields + the -8 corner cases
        flat = v.reshape(-1)
        flat[:8] = [0xFF, 0x88, 0x80, 0x08, 0x00, 0x77, 0x78, 0x87]
        flat[-4:] = [0x88, 0x88, 0xFF, 0x80]
    return v


def _run_corr(torch, v, impl=0, ngulp=1):
    from xengine import ops
    nt, nchan, nstand, npol = v.shape
    nt_g = nt // ngulp
    xc = ops.XCorr(nstand=nstand, npol=npol, nchan=nchan, ntime_gulp=nt_g)
    xc.set_impl(impl)
    out = torch.full((2, nchan, ops.words_per_chan(nstand)), -12345, dtype=torch.int32, device="cuda")
    for g in range(ngulp):
        d_in = torch.from_numpy(np.ascontiguousarray(v[g * nt_g:(g + 1) * nt_g])).cuda()
        xc.run(d_in, out, g == ngulp - 1)
        xc.sync()
    res = out.cpu().numpy()
    xc.close()
    return res


@pytest.mark.parametrize("impl", [1, 2, 0])   # CUDA-core check, register-fed tcgen05, TMA-fed tcgen05 (default)
@pytest.mark.parametrize("shape", [
    (16, 2, 4), (37, 3, 8), (32, 1, 64), (50, 3, 68), (480, 2, 352), (96, 5, 128), (17, 2, 12),
])
def test_corr_matches_oracle(torch_cuda, impl, shape):
    nt, nchan, nstand = shape
    rng = np.random.default_rng(1234 + nt + nstand)
    v = _voltages(rng, nt, nchan, nstand)
    got = _run_corr(torch_cuda, v, impl=impl)
    want = orc.corr_xgpu(v)
    assert got.dtype == np.int32 and got.shape == want.shape
    assert np.array_equal(got, want)


def test_corr_extreme_values(torch_cuda):
    """All inputs -8-8j: every product is 128, the largest magnitude; and all-zero input."""
    v = np.full((480, 2, 352, 2), 0x88, dtype=np.uint8)
    got = _run_corr(torch_cuda, v)
    want = orc.corr_xgpu(v)
    assert np.array_equal(got, want) and got.max() == 128 * 480
    v[:] = 0
    assert not _run_corr(torch_cuda, v).any()


@pytest.mark.parametrize("impl", [1, 2, 0])   # CUDA-core check, register-fed tcgen05, TMA-fed tcgen05 (default)
def test_corr_multi_gulp_dump_and_clear(torch_cuda, impl):
    """xGPU semantics (corr_block.py:445): integrate over calls, write on doDump, clear afterwards."""
    from xengine import ops
    torch = torch_cuda
    rng = np.random.default_rng(7)
    nt, nchan, nstand = 48, 3, 36
    xc = ops.XCorr(nstand=nstand, nchan=nchan, ntime_gulp=nt)
    xc.set_impl(impl)
    out = torch.zeros((2, nchan, ops.words_per_chan(nstand)), dtype=torch.int32, device="cuda")
    for integ, ngulp in enumerate([3, 1, 2]):
        want = np.zeros(out.shape, dtype=np.int64)
        out.fill_(-1)
        for g in range(ngulp):
            v = _voltages(rng, nt, nchan, nstand)
            want += orc.corr_xgpu(v)
            xc.run(torch.from_numpy(v).cuda(), out, g == ngulp - 1)
            xc.sync()
            if g < ngulp - 1:
                assert int(out.min()) == -1 and int(out.max()) == -1   # untouched until doDump
        assert np.array_equal(out.cpu().numpy(), want.astype(np.int32)), "integration %d" % integ
    xc.close()


def test_corr_fused_accumulate(torch_cuda):
    """Fused Corr+CorrAcc (bulk-reduce epilogue) == oracle corr followed by corr_acc."""
    from xengine import ops
    torch = torch_cuda
    rng = np.random.default_rng(11)
    nt, nchan, nstand = 480, 2, 352
    xc = ops.XCorr(nstand=nstand, nchan=nchan, ntime_gulp=nt)
    acc = torch.full((2, nchan, ops.words_per_chan(nstand)), 99, dtype=torch.int32, device="cuda")
    want = None
    for g in range(3):
        v = _voltages(rng, nt, nchan, nstand)
        x = orc.corr_xgpu(v)
        want = orc.corr_acc(x.copy() if want is None else want, x, first=(g == 0))
        xc.accumulate(torch.from_numpy(v).cuda(), acc, g == 0)
    xc.sync()
    assert np.array_equal(acc.cpu().numpy(), want)
    xc.close()


def test_corr_host_variant(torch_cuda):
    """bfXgpuCorrelate: host buffers in and out (xgpu_test.py:86-89)."""
    from xengine import ops
    rng = np.random.default_rng(5)
    v = _voltages(rng, 64, 2, 16)
    xc = ops.XCorr(nstand=16, nchan=2, ntime_gulp=64)
    out = np.zeros((2, 2, ops.words_per_chan(16)), dtype=np.int32)
    xc.run_host(v, out, True)
    assert np.array_equal(out, orc.corr_xgpu(v))
    xc.close()


def test_corr_tc_equals_simt_full_size(torch_cuda):
    """704 inputs x 480 samples, more channels than the oracle is asked to do: tensor-core kernel vs the
    CUDA-core cross-check kernel on every word, oracle on two channels."""
    rng = np.random.default_rng(2024)
    v = _voltages(rng, 480, 24, 352)
    a = _run_corr(torch_cuda, v, impl=0)
    b = _run_corr(torch_cuda, v, impl=1)
    assert np.array_equal(a, b)
    for c in (0, 23):
        assert np.array_equal(a[:, c:c + 1], orc.corr_xgpu(v[:, c:c + 1]))


def test_corr_linearity_in_time(torch_cuda):
    """Size-independent property: X(gulp A ++ gulp B) == X(A) + X(B)."""
    rng = np.random.default_rng(99)
    v = _voltages(rng, 960, 4, 352)
    whole = _run_corr(torch_cuda, v, ngulp=1)
    two = _run_corr(torch_cuda, v, ngulp=2)
    parts = _run_corr(torch_cuda, v[:480]).astype(np.int64) + _run_corr(torch_cuda, v[480:])
    assert np.array_equal(whole, two) and np.array_equal(whole, parts.astype(np.int32))


def test_get_order_matches_oracle():
    from xengine import ops
    for nstand in (4, 12, 352):
        a2i = np.arange(nstand * 2, dtype=np.int32).reshape(nstand, 2)
        for perm in (False, True):
            if perm:
                a2i = np.random.default_rng(3).permutation(nstand * 2).astype(np.int32).reshape(nstand, 2)
            a2bl, conj = ops.get_order(a2i)
            wa, wc = orc.get_order(a2i)
            assert np.array_equal(a2bl, wa) and np.array_equal(conj, wc)


@pytest.mark.parametrize("n", [0, 1, 5, 4096, 1000003])
def test_acc(torch_cuda, n):
    from xengine import ops
    torch = torch_cuda
    rng = np.random.default_rng(n)
    a = rng.integers(-2**30, 2**30, size=n, dtype=np.int32)
    b = rng.integers(-2**30, 2**30, size=n, dtype=np.int32)
    da, db = torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda()
    if n == 0:
        da = torch.zeros(4, dtype=torch.int32, device="cuda"); db = da.clone()
    ops.acc(da, db, n, True)
    torch.cuda.synchronize()
    assert np.array_equal(da.cpu().numpy()[:n], b)
    ops.acc(da, db, n, False)
    ops.acc(da, db, n, False)
    torch.cuda.synchronize()
    assert np.array_equal(da.cpu().numpy()[:n], (3 * b.astype(np.int64)).astype(np.int32))


@pytest.mark.parametrize("nstand,nchan", [(12, 8), (72, 40), (132, 33), (352, 5)])
def test_reorder_tiled_and_packet_payloads(torch_cuda, nstand, nchan):
    """The tiled reorder (full square matrix) and the per-baseline packet payloads against the oracle, on random
    buffer contents (the kernel is a pure permutation, so it does not need a real correlation), with a shuffled
    antenna map for one case so nothing depends on the identity map."""
    from xengine import ops
    torch = torch_cuda
    rng = np.random.default_rng(nstand * 1000 + nchan)
    wpc = ops.words_per_chan(nstand)
    x = rng.integers(-2**31, 2**31 - 1, size=(2, nchan, wpc), dtype=np.int64).astype(np.int32)
    a2i = np.arange(nstand * 2, dtype=np.int32).reshape(nstand, 2)
    if nstand == 72:
        a2i = rng.permutation(nstand * 2).astype(np.int32).reshape(nstand, 2)
    a2bl, conj = ops.get_order(a2i)
    dx, da, dc = torch.from_numpy(x).cuda(), torch.from_numpy(a2bl).cuda(), torch.from_numpy(conj).cuda()
    full = torch.full((nstand, nstand, 2, 2, nchan, 2), 0x5a5a5a5a, dtype=torch.int32, device="cuda")
    ops.reorder(dx, full, da, dc, nstand, 2, nchan, wpc)
    want = orc.reorder(x, a2bl, conj)
    assert np.array_equal(full.cpu().numpy(), want)
    nbl = nstand * (nstand + 1) // 2
    for layout, name in ((ops.PAYLOAD_COR, "cor"), (ops.PAYLOAD_PY, "py")):
        out = torch.full((nbl * nchan * 8,), 0x5a5a5a5a, dtype=torch.int32, device="cuda")
        ops.reorder_packets(dx, out, da, dc, nstand, 2, nchan, wpc, layout)
        exp = orc.packet_payloads(want, name)
        assert np.array_equal(out.cpu().numpy().reshape(exp.shape), exp), name
    with pytest.raises(Exception):
        ops.reorder_packets(dx, out, da, dc, nstand, 1, nchan, wpc, ops.PAYLOAD_COR)


def test_subsel_and_reorder(torch_cuda):
    from xengine import ops
    torch = torch_cuda
    rng = np.random.default_rng(42)
    nt, nchan, nstand = 40, 8, 12
    v = _voltages(rng, nt, nchan, nstand)
    x = _run_corr(torch, v)
    assert np.array_equal(x, orc.corr_xgpu(v))
    a2i = np.arange(nstand * 2, dtype=np.int32).reshape(nstand, 2)
    a2bl, conj = ops.get_order(a2i)
    # reorder == v(s0,p0) conj(v(s1,p1)) for every pair, which is what make_golden_inputs.py pins
    dx = torch.from_numpy(x).cuda()
    out = torch.zeros((nstand, nstand, 2, 2, nchan, 2), dtype=torch.int32, device="cuda")
    ops.reorder(dx, out, torch.from_numpy(a2bl).cuda(), torch.from_numpy(conj).cuda(), nstand, 2, nchan,
                ops.words_per_chan(nstand))
    torch.cuda.synchronize()
    got = out.cpu().numpy()
    assert np.array_equal(got, orc.reorder(x, a2bl, conj))
    fre, fim = orc.corr_full(v)
    assert np.array_equal(got[..., 0], np.moveaxis(fre, 0, -1))
    assert np.array_equal(got[..., 1], np.moveaxis(fim, 0, -1))
    # subselect: random baselines in both station orders, nchan_sum in {1,2,4}
    bls = [[[int(rng.integers(nstand)), int(rng.integers(2))], [int(rng.integers(nstand)), int(rng.integers(2))]]
           for _ in range(301)]
    vm, cj = orc.baselines_to_vismap(bls, a2bl, conj)
    for ncs in (1, 2, 4):
        o = torch.zeros((nchan // ncs, len(bls), 2), dtype=torch.int32, device="cuda")
        ops.subsel(dx, o, torch.from_numpy(vm).cuda(), torch.from_numpy(cj).cuda(), len(bls), nchan, ncs,
                   ops.words_per_chan(nstand))
        torch.cuda.synchronize()
        got = o.cpu().numpy()
        assert np.array_equal(got, orc.subselect(x, vm, cj, ncs))
        # test_corr_part_rx.py:76-84: equals the channel-summed logical visibility
        for k, ((s0, p0), (s1, p1)) in enumerate(bls[:50]):
            assert np.array_equal(got[:, k, 0], fre[:, s0, s1, p0, p1].reshape(-1, ncs).sum(1))
            assert np.array_equal(got[:, k, 1], fim[:, s0, s1, p0, p1].reshape(-1, ncs).sum(1))
    # the same selection written in CorrOutputPart's packet orders (corr_output_part_block.py:378-391, 429-463)
    nv = 300
    want = orc.subselect(x, vm[:nv], cj[:nv], 2)
    dvm, dcj = torch.from_numpy(vm[:nv].copy()).cuda(), torch.from_numpy(cj[:nv].copy()).cuda()
    for layout, name in ((ops.PAYLOAD_PY, "py"), (ops.PAYLOAD_COR, "cor")):
        o = torch.zeros(nchan // 2 * nv * 2, dtype=torch.int32, device="cuda")
        ops.subsel_packets(dx, o, dvm, dcj, nv, nchan, 2, ops.words_per_chan(nstand), layout)
        exp = orc.subsel_packet_payloads(want, name)
        assert o.cpu().numpy().tobytes() == exp.tobytes(), name
    with pytest.raises(Exception):
        ops.subsel_packets(dx, o, dvm, dcj, 301, nchan, 2, ops.words_per_chan(nstand), ops.PAYLOAD_COR)


@pytest.mark.parametrize("shape,ncs,impl", [
    ((96, 8, 64), 4, 0),     # TMA kernel, grouped walk (channel sum inside a CTA), lists longer than 64 entries
    ((96, 8, 64), 1, 0),     # nchan_sum = 1: every tile stores its own rows
    ((64, 16, 100), 8, 0),   # groups of 8 channels, 10 tile positions
    ((96, 32, 64), 16, 0),   # nchan_sum > 8: ungrouped walk, zeroing + reductions across CTAs
    ((130, 4, 68), 2, 0),    # ragged last row block, K not a multiple of the stage
    ((40, 8, 12), 4, 0),     # ninput = 24: not a multiple of 16 -> unfused fallback inside the library
    ((64, 4, 36), 1, 2),     # register-fed kernel -> unfused fallback
])
def test_corr_fused_accumulate_subsel(torch_cuda, shape, ncs, impl):
    """BASELINE configs[4]: Corr + CorrAcc + CorrSubsel in one call; accumulator and fast dump both bit-exact."""
    from xengine import ops
    torch = torch_cuda
    nt, nchan, nstand = shape
    rng = np.random.default_rng(99 + nstand)
    a2bl, conj = ops.get_order(np.arange(nstand * 2, dtype=np.int32).reshape(nstand, 2))
    # default selection of the reference (pol-0 autos, corr_subsel_block.py:230) + random baselines in both orders
    bls = [[[i % nstand, 0], [i % nstand, 0]] for i in range(48)]
    bls += [[[int(rng.integers(nstand)), int(rng.integers(2))], [int(rng.integers(nstand)), int(rng.integers(2))]]
            for _ in range(700)]
    vm, cj = orc.baselines_to_vismap(bls, a2bl, conj)
    xc = ops.XCorr(nstand=nstand, nchan=nchan, ntime_gulp=nt)
    xc.set_impl(impl)
    xc.set_subsel(vm, cj, ncs)
    wpc = ops.words_per_chan(nstand)
    acc = torch.full((2, nchan, wpc), 7, dtype=torch.int32, device="cuda")
    fast = torch.full((nchan // ncs, len(bls), 2), -5, dtype=torch.int32, device="cuda")
    want_acc = np.zeros((2, nchan, wpc), dtype=np.int64)
    for g in range(3):
        v = _voltages(rng, nt, nchan, nstand)
        x = orc.corr_xgpu(v)
        want_acc += x
        xc.accumulate_subsel(torch.from_numpy(v).cuda(), acc, g == 0, fast)
        xc.sync()
        assert np.array_equal(fast.cpu().numpy(), orc.subselect(x, vm, cj, ncs)), "fast dump of gulp %d" % g
        assert np.array_equal(acc.cpu().numpy(), want_acc.astype(np.int32)), "accumulator after gulp %d" % g
    # sub_first = False: the fast dump integrates over several gulps (Corr's acc_len > ntime_gulp)
    fast.fill_(-5)
    want_fast = np.zeros((nchan // ncs, len(bls), 2), dtype=np.int64)
    for g in range(3):
        v = _voltages(rng, nt, nchan, nstand)
        want_fast += orc.subselect(orc.corr_xgpu(v), vm, cj, ncs)
        xc.accumulate_subsel(torch.from_numpy(v).cuda(), acc, False, fast, g == 0)
    xc.sync()
    assert np.array_equal(fast.cpu().numpy(), want_fast.astype(np.int32)), "fast dump integrated over 3 gulps"
    xc.set_subsel(np.zeros(0, np.int32), np.zeros(0, np.int32), ncs)     # cleared: the call must refuse
    with pytest.raises(Exception):
        xc.accumulate_subsel(torch.from_numpy(v).cuda(), acc, False, fast)
    xc.close()


def test_corr_fused_subsel_full_size(torch_cuda):
    """704 inputs x 192 channels x 480 samples, 4704 and 4656 baselines, nchan_sum = 4: the fused fast dump equals the
    stand-alone CorrSubsel kernel run on the stand-alone Corr output (device-side cross-check at BASELINE size)."""
    from xengine import ops
    torch = torch_cuda
    nt, nchan, nstand = 480, 192, 352
    rng = np.random.default_rng(5)
    a2bl, conj = ops.get_order(np.arange(nstand * 2, dtype=np.int32).reshape(nstand, 2))
    wpc = ops.words_per_chan(nstand)
    v = torch.randint(0, 256, (nt, nchan, nstand * 2), dtype=torch.uint8, device="cuda")
    xc = ops.XCorr(nstand=nstand, nchan=nchan, ntime_gulp=nt)
    ref = torch.zeros((2, nchan, wpc), dtype=torch.int32, device="cuda")
    xc.run(v, ref, True)
    xc.sync()
    for nvis in (4704, 4656):
        sel = rng.integers(0, a2bl.size, nvis)
        vm, cj = a2bl.reshape(-1)[sel].copy(), conj.reshape(-1)[sel].copy()
        xc.set_subsel(vm, cj, 4)
        acc = torch.zeros((2, nchan, wpc), dtype=torch.int32, device="cuda")
        fast = torch.zeros((nchan // 4, nvis, 2), dtype=torch.int32, device="cuda")
        xc.accumulate_subsel(v, acc, True, fast)
        xc.accumulate_subsel(v, acc, False, fast)
        xc.sync()
        want = torch.zeros_like(fast)
        ops.subsel(ref, want, torch.from_numpy(vm).cuda(), torch.from_numpy(cj).cuda(), nvis, nchan, 4, wpc)
        torch.cuda.synchronize()
        assert torch.equal(fast, want)
        assert torch.equal(acc, 2 * ref)
        del acc, fast
    xc.close()


def test_corr_baseline_size_properties(torch_cuda):
    """BASELINE configs[2] per-GPU size (704 inputs x 384 channels x 480 samples): size-independent properties checked on
    the device over the whole 766 MB matrix, the oracle on three spot channels.
      * autocorrelation words: imaginary part 0, real part = sum_t |v|^2 (every channel, every input);
      * same-stand cross-pol pair: word(in0=2s, in1=2s+1) == conj-symmetric partner values via the oracle spot check;
      * linearity: accumulating the gulp twice gives exactly 2x the single-gulp dump."""
    from xengine import ops
    torch = torch_cuda
    nt, nchan, nstand = 480, 384, 352
    ninput = 2 * nstand
    wpc = ops.words_per_chan(nstand)
    g = torch.Generator(device="cuda")
    g.manual_seed(0xdeadbeef)
    v = torch.randint(0, 256, (nt, nchan, ninput), dtype=torch.uint8, device="cuda", generator=g)
    xc = ops.XCorr(nstand=nstand, nchan=nchan, ntime_gulp=nt)
    once = torch.zeros((2, nchan, wpc), dtype=torch.int32, device="cuda")
    xc.run(v, once, True)
    twice = torch.full((2, nchan, wpc), -1, dtype=torch.int32, device="cuda")
    xc.accumulate(v, twice, True)
    xc.accumulate(v, twice, False)
    xc.sync()
    assert torch.equal(twice, 2 * once)
    del twice
    # autocorrelations
    a2bl, _ = ops.get_order(np.arange(ninput, dtype=np.int32).reshape(nstand, 2))
    auto_idx = torch.from_numpy(np.array([a2bl[s, s, p, p] for s in range(nstand) for p in range(2)], dtype=np.int64)).cuda()
    vi = v.to(torch.int32)
    re = (vi >> 4) - ((vi >> 7) << 4)                  # sign-extend the high nibble
    im = (vi & 15) - (((vi >> 3) & 1) << 4)            # and the low one
    power = (re * re + im * im).sum(dim=0)             # [nchan][ninput]
    assert torch.equal(once[0][:, auto_idx], power)
    assert not once[1][:, auto_idx].any()
    # oracle on three channels
    got = once.cpu().numpy()
    vh = v.cpu().numpy().reshape(nt, nchan, nstand, 2)
    for c in (0, 191, 383):
        assert np.array_equal(got[:, c:c + 1], orc.corr_xgpu(vh[:, c:c + 1]))
    xc.close()


def test_reorder_arbitrary_tables_take_the_general_path(torch_cuda):
    """xeng_reorder is table-driven: with a2bl / conj tables that are NOT what get_order builds (random word indices,
    so a mirrored entry is not one of the four words of its stand pair) the kernel must fall back from the
    tile-reuse shortcut to gathering every entry, and still equal the oracle's plain gather."""
    from xengine import ops
    torch = torch_cuda
    nstand, nchan = 40, 37
    rng = np.random.default_rng(99)
    wpc = ops.words_per_chan(nstand)
    x = rng.integers(-2**31, 2**31 - 1, size=(2, nchan, wpc), dtype=np.int64).astype(np.int32)
    a2bl = rng.integers(0, wpc, size=(nstand, nstand, 2, 2)).astype(np.int32)
    conj = rng.integers(0, 2, size=(nstand, nstand, 2, 2)).astype(np.int32)
    full = torch.full((nstand, nstand, 2, 2, nchan, 2), 0x5a5a5a5a, dtype=torch.int32, device="cuda")
    ops.reorder(torch.from_numpy(x).cuda(), full, torch.from_numpy(a2bl).cuda(), torch.from_numpy(conj).cuda(),
                nstand, 2, nchan, wpc)
    assert np.array_equal(full.cpu().numpy(), orc.reorder(x, a2bl, conj))


def test_config2_full_dump_every_channel_with_fused_fast_dump(torch_cuda):
    """BASELINE configs[1]/[4] at full width, EVERY channel against the oracle ("bit-exact vs
    oracle in xGPU order for >= 1 full dump at reduced length"): 704 inputs x 192 channels x 480-sample gulps,
    10 spans accumulated by the fused Corr+CorrAcc+CorrSubsel launch.  Per span the fast dump (4704 baselines for the
    first five spans, 4656 for the rest, nchan_sum = 4) equals orc.subselect(orc.corr_xgpu(span)); the long
    accumulator equals the sum of the ten oracle matrices on all 2 x 192 x 249 216 words.  Voltages as the reference
    generates them (make_golden_inputs.py:57,137: RandomState(0xdeadbeef).randint(0, 255)) plus the edge bytes that
    randint(0, 255) never yields.  The persistent tile walk wraps channels across the 148 CTAs: only a run at the full
    channel count visits every wrap."""
    from xengine import ops
    torch = torch_cuda
    nt, nchan, nstand, nspan = 480, 192, 352, 10
    wpc = ops.words_per_chan(nstand)
    rs = np.random.RandomState(0xdeadbeef)
    rng = np.random.default_rng(17)
    a2bl, conj = ops.get_order(np.arange(nstand * 2, dtype=np.int32).reshape(nstand, 2))
    # 4704 = 48 stands x 48 stands lower triangle x 4 pol products (corr_subsel_block.py:185), both station orders;
    # 4656 (docs/source/outputs.rst:369) = random words
    st = rng.choice(nstand, 48, replace=False)
    bls_a = [[[int(s0), p0], [int(s1), p1]] for i, s0 in enumerate(st) for s1 in st[:i + 1]
             for p0 in range(2) for p1 in range(2)]
    assert len(bls_a) == 4704
    bls_b = [[[int(rng.integers(nstand)), int(rng.integers(2))], [int(rng.integers(nstand)), int(rng.integers(2))]]
             for _ in range(4656)]
    sel = {4704: orc.baselines_to_vismap(bls_a, a2bl, conj), 4656: orc.baselines_to_vismap(bls_b, a2bl, conj)}
    xc = ops.XCorr(nstand=nstand, nchan=nchan, ntime_gulp=nt)
    acc = torch.full((2, nchan, wpc), 3, dtype=torch.int32, device="cuda")
    want_acc = np.zeros((2, nchan, wpc), dtype=np.int64)
    for g in range(nspan):
        nvis = 4704 if g < 5 else 4656
        vm, cj = sel[nvis]
        if g in (0, 5):
            xc.set_subsel(vm, cj, 4)
            fast = torch.full((nchan // 4, nvis, 2), -9, dtype=torch.int32, device="cuda")
        v = rs.randint(0, 255, size=(nt, nchan, nstand, 2), dtype=np.uint8)
        flat = v.reshape(-1)
        flat[:8] = [0xFF, 0x88, 0x80, 0x08, 0x00, 0x77, 0x78, 0x87]
        flat[-4:] = [0x88, 0x88, 0xFF, 0x80]
        xc.accumulate_subsel(torch.from_numpy(v).cuda(), acc, g == 0, fast)      # runs while the oracle computes
        x = orc.corr_xgpu(v)
        want_acc += x
        xc.sync()
        assert np.array_equal(fast.cpu().numpy(), orc.subselect(x, vm, cj, 4)), "fast dump of span %d" % g
    got = acc.cpu().numpy()
    assert np.array_equal(got, want_acc.astype(np.int32))
    assert np.abs(want_acc).max() < 2 ** 31
    xc.close()
