"""
CPU oracle for the LWA-352 X-engine hot path.  TEST INFRASTRUCTURE ONLY.

This module restates, in numpy, the arithmetic that the reference
(realtimeradio/caltech-bifrost-dsp) pins in its own verification code.  It is
the checker for the CUDA path; it is never the product.  Only ``tests/``,
``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference``
legs of ``bench.py`` may import it.

Parity status: PINNED.  The algorithm lives in two git submodules that are
absent from /root/reference (``bifrost`` @ 474d87016b9a, ``xgpu`` @ cb5dba40d46d),
so the oracle follows the reference's own numpy restatements and is pinned
against fixtures produced by *executing those reference functions*
(``tests/golden/make_golden.py`` imports them from /root/reference with the
absent ``bifrost``/``ujson``/``etcd3`` modules stubbed) -- see
``tests/test_oracle_golden.py``.  Exception: the offline upchannelise functions at
the end restate stock bifrost blocks for which the reference holds
no test or vector -- for those, PARITY UNPINNED by reference vectors; they are
checked against the literal DFT definition only.

Every function cites the reference file:line it follows (paths relative to
/root/reference).
"""
import numpy as np

# ---------------------------------------------------------------------------
# 4+4-bit complex decode
# ---------------------------------------------------------------------------

def unpack4(u8):
    """byte = (re << 4) | (im & 0xf), both 4-bit two's complement.

    Follows pipeline/lwa352_pipeline/blocks/corr_block.py:270-275
    (identical code at pipeline/verification/make_golden_inputs.py:117-121 and
    pipeline/verification/tests/beamformer/beamformer_test.py:68-73).
    Returns (re, im) as int8 arrays of the input's shape.
    """
    d = np.asarray(u8).view(np.uint8) if np.asarray(u8).dtype != np.uint8 else np.asarray(u8)
    dr = (d >> 4).astype(np.int8)
    dr[dr > 7] -= 16
    di = (d & 0xF).astype(np.int8)
    di[di > 7] -= 16
    return dr, di


# ---------------------------------------------------------------------------
# xGPU register-tile order
# ---------------------------------------------------------------------------

def tri_index(i, j):
    """corr_block.py:27-28 (valid for i >= j)."""
    return (i * (i + 1)) // 2 + j


def regtile_index(in0, in1, nstand):
    """Word index of the real part of conj(v[in0]) * v[in1] inside one channel
    of the xGPU register-tile-ordered buffer; valid when station(in1) >=
    station(in0).  corr_block.py:37-58 (dup pipeline/verification/xgpu_test.py:31-52).
    Works on scalars or integer arrays.
    """
    a0 = in0 >> 1
    a1 = in1 >> 1
    p0 = in0 & 1
    p1 = in1 & 1
    quadrant_index = tri_index(a1 // 2, a0 // 2)
    quadrant = 2 * (a0 & 1) + (a1 & 1)
    quadrant_size = (nstand // 2 + 1) * nstand // 4
    cell_index = quadrant * quadrant_size + quadrant_index
    return cell_index * 4 + 2 * p1 + p0


def words_per_chan(nstand, npol=2):
    """Per-channel length of the xGPU buffer: corr_block.py:231 with nchan=1."""
    return (nstand // 2 + 1) * (nstand // 4) * npol * npol * 4


# ---------------------------------------------------------------------------
# Correlator
# ---------------------------------------------------------------------------

def _gram(re, im):
    """Exact G[i,j] = sum_t conj(v[t,i]) * v[t,j] for one channel, via float
    matmuls that are exact for these magnitudes (|term| <= 128, so float32 is
    exact for ntime < 2**24/128; float64 otherwise).
    Returns (Gre, Gim) as int64 [ninput, ninput].
    """
    nt = re.shape[0]
    ft = np.float32 if nt * 128 < (1 << 24) else np.float64
    R = re.astype(ft)
    I = im.astype(ft)
    rr = R.T @ R
    ii = I.T @ I
    ri = R.T @ I          # sum_t r_i * i_j
    gre = rr + ii
    gim = ri - ri.T       # r_i i_j - i_i r_j
    return np.rint(gre).astype(np.int64), np.rint(gim).astype(np.int64)


def corr_full(v_u8):
    """Logical-order correlation of one integration.

    v_u8: uint8 [ntime, nchan, nstand, npol].
    Returns (re, im) int64 [nchan, nstand, nstand, npol, npol] with
    out[c,s0,s1,p0,p1] = sum_t v(s0,p0) * conj(v(s1,p1)),
    the convention of pipeline/verification/make_golden_inputs.py:149-158.
    """
    nt, nchan, nstand, npol = v_u8.shape
    re, im = unpack4(v_u8)
    ni = nstand * npol
    ore = np.empty([nchan, nstand, nstand, npol, npol], dtype=np.int64)
    oim = np.empty_like(ore)
    for c in range(nchan):
        gre, gim = _gram(re[:, c].reshape(nt, ni), im[:, c].reshape(nt, ni))
        # G[i,j] = conj(v_i) v_j  ->  v_i conj(v_j) = conj(G[i,j])
        gre = gre.reshape(nstand, npol, nstand, npol).transpose(0, 2, 1, 3)
        gim = gim.reshape(nstand, npol, nstand, npol).transpose(0, 2, 1, 3)
        ore[c] = gre
        oim[c] = -gim
    return ore, oim


def corr_loop_literal(v_u8):
    """Literal restatement of Corr._test (corr_block.py:265-287): a Python loop
    of np.outer over time and channel.  Very slow; small cases and the timed
    CPU baseline only.  Returns complex128 [nchan, nstand, nstand, npol*npol]
    with out[c,i,j,2*p0+p1] = sum_t conj(v(i,p0)) * v(j,p1).
    """
    nt, nchan, nstand, npol = v_u8.shape
    dr, di = unpack4(v_u8)
    dc = dr + 1j * di
    out = np.zeros([nchan, nstand, nstand, npol * npol], dtype=complex)
    for t in range(nt):
        for chan in range(nchan):
            out[chan, :, :, 0] += np.outer(np.conj(dc[t, chan, :, 0]), dc[t, chan, :, 0])
            out[chan, :, :, 1] += np.outer(np.conj(dc[t, chan, :, 0]), dc[t, chan, :, 1])
            out[chan, :, :, 2] += np.outer(np.conj(dc[t, chan, :, 1]), dc[t, chan, :, 0])
            out[chan, :, :, 3] += np.outer(np.conj(dc[t, chan, :, 1]), dc[t, chan, :, 1])
    return out


_XGPU_MAP_CACHE = {}


def xgpu_maps(nstand, npol=2):
    """For every (in0,in1) with station(in1) >= station(in0): the word index
    in the per-channel xGPU buffer.  Returns (in0, in1, idx) int64 arrays."""
    key = (nstand, npol)
    if key not in _XGPU_MAP_CACHE:
        assert npol == 2, "xGPU register-tile order is defined for npol == 2"
        ni = nstand * npol
        in0, in1 = np.meshgrid(np.arange(ni), np.arange(ni), indexing="ij")
        ok = (in1 >> 1) >= (in0 >> 1)
        in0 = in0[ok]
        in1 = in1[ok]
        _XGPU_MAP_CACHE[key] = (in0, in1, regtile_index(in0, in1, nstand))
    return _XGPU_MAP_CACHE[key]


def corr_xgpu(v_u8, acc=None):
    """Raw Corr output: int32 [2 (re,im), nchan, W] in xGPU register-tile order.

    Word regtile_index(in0,in1) (station(in1) >= station(in0)) holds
    sum_t conj(v[in0]) * v[in1]  -- pipeline/verification/xgpu_test.py:102-136
    and corr_block.py:289-306; imaginary plane is ``matlen`` words after the
    real plane (corr_block.py:292-293).  The nstand/2*4 padding words per
    channel (quadrant 2, diagonal cells) are never read by any reference test;
    this oracle (and the CUDA path) leave them 0.
    If ``acc`` is given the result is added to it (xGPU accumulates over calls
    until doDump, corr_block.py:445).
    """
    nt, nchan, nstand, npol = v_u8.shape
    W = words_per_chan(nstand, npol)
    re, im = unpack4(v_u8)
    ni = nstand * npol
    in0, in1, idx = xgpu_maps(nstand, npol)
    out = np.zeros([2, nchan, W], dtype=np.int64) if acc is None else acc
    for c in range(nchan):
        gre, gim = _gram(re[:, c].reshape(nt, ni), im[:, c].reshape(nt, ni))
        out[0, c, idx] += gre[in0, in1]
        out[1, c, idx] += gim[in0, in1]
    if acc is None:
        assert np.abs(out).max() < 2 ** 31
        return out.astype(np.int32)
    return out


def corr_acc(acc, x, first):
    """CorrAcc vector accumulator: ``a = b`` on the first span of an
    integration, ``a += b`` afterwards (corr_acc_block.py:299-306)."""
    if first:
        acc[...] = x
    else:
        acc += x
    return acc


# ---------------------------------------------------------------------------
# Baseline order maps, subselect, reorder
# ---------------------------------------------------------------------------

def get_order(ant_to_input, nstand=None, npol=None):
    """bfXgpuGetOrder (call site corr_block.py:317-333).

    For every (s0,s1,p0,p1) returns the word index in the per-channel xGPU
    buffer that holds that pair, and whether that word must be conjugated to
    read as v(s0,p0) * conj(v(s1,p1)) -- the net convention pinned by
    corr_output_full_block.py:582-591 and
    pipeline/verification/test_corr_part_rx.py:76-84.
    The xGPU word (in0,in1) holds conj(v[in0]) v[in1], so when the pair is
    stored as (i0,i1) it is the conjugate of what is wanted (conj=1); when it
    has to be looked up swapped, (i1,i0), it is already right (conj=0).
    Source of bfXgpuGetOrder itself is absent (bifrost submodule).
    """
    a2i = np.asarray(ant_to_input, dtype=np.int64)
    nstand, npol = a2i.shape
    i0 = a2i[:, None, :, None]       # [s0, 1, p0, 1]
    i1 = a2i[None, :, None, :]       # [1, s1, 1, p1]
    i0, i1 = np.broadcast_arrays(i0, i1)
    stored = (i1 >> 1) >= (i0 >> 1)
    lo = np.where(stored, i0, i1)
    hi = np.where(stored, i1, i0)
    a2bl = regtile_index(lo, hi, nstand).astype(np.int32)
    conj = stored.astype(np.int32)
    # same-station pairs are stored both ways round; either is valid.  Use the
    # direct (i0,i1) slot, conj=1, as `stored` already says.
    return a2bl, conj


def subselect(xbuf, vismap, conj, nchan_sum):
    """bfXgpuSubSelect (call site corr_subsel_block.py:298; semantics pinned by
    pipeline/verification/test_corr_part_rx.py:38-49,76-84).

    xbuf: int32 [2, nchan, W]; vismap/conj: int32 [nvis].
    Returns int32 [nchan//nchan_sum, nvis, 2] (re, im interleaved:
    corr_subsel_block.py:302, corr_output_part_block.py:455) with
    out[c',v] = sum_k xbuf[:, c'*nchan_sum + k, vismap[v]], imag negated
    where conj[v] != 0.
    """
    _, nchan, W = xbuf.shape
    nco = nchan // nchan_sum
    g = xbuf[:, : nco * nchan_sum, :][:, :, np.asarray(vismap)].astype(np.int64)
    g = g.reshape(2, nco, nchan_sum, -1).sum(axis=2)
    sign = np.where(np.asarray(conj) != 0, -1, 1)
    out = np.empty([nco, g.shape[-1], 2], dtype=np.int32)
    out[..., 0] = g[0]
    out[..., 1] = g[1] * sign
    return out


def baselines_to_vismap(baselines, a2bl, conj):
    """CorrSubsel.update_subsel (corr_subsel_block.py:237-250)."""
    vm = np.empty(len(baselines), dtype=np.int32)
    cj = np.empty(len(baselines), dtype=np.int32)
    for v, ((s0, p0), (s1, p1)) in enumerate(baselines):
        vm[v] = a2bl[s0, s1, p0, p1]
        cj[v] = conj[s0, s1, p0, p1]
    return vm, cj


def reorder(xbuf, a2bl, conj):
    """bfXgpuReorder (call site corr_output_full_block.py:669, output shape
    :392, convention :582-591).  Returns int32
    [nstand, nstand, npol, npol, nchan, 2] = v(s0,p0) conj(v(s1,p1))."""
    g = xbuf[:, :, a2bl]                         # [2, nchan, s0, s1, p0, p1]
    sign = np.where(conj != 0, -1, 1)
    out = np.empty(a2bl.shape + (xbuf.shape[1], 2), dtype=np.int32)
    out[..., 0] = np.moveaxis(g[0], 0, -1)
    out[..., 1] = np.moveaxis(g[1], 0, -1) * sign[..., None]
    return out


def snap2_packet(seq, sync_time, npol, npol_tot, nchan, nchan_tot, chan_block_id, chan0, pol0, payload):
    """One SNAP2 F-engine packet as the reference's transmitters build it: struct.pack('>QLHHHHLLL', ...) + payload
    u8 [nchan][npol] (pipeline/test_transmitters/test_tx_vectors.py:37-47,103-109; test_tx.c:27-42)."""
    import struct
    return struct.pack('>QLHHHHLLL', seq, sync_time, npol, npol_tot, nchan, nchan_tot, chan_block_id, chan0, pol0) + \
        np.ascontiguousarray(payload, dtype=np.uint8).tobytes()


def snap2_assemble(packets, seq0, ntime, nchan, ninput, first_chan=0):
    """What the capture leaves in the gpu-input ring for a window of spectra: out[seq - seq0][chan0 - first_chan + c]
    [pol0 + p] = payload[c][p], zeros where no packet arrived, packets outside the window dropped
    (capture_block.py:219-305 + the header above).  Returns (u8 [ntime, nchan, ninput], accepted packet count)."""
    import struct
    out = np.zeros((ntime, nchan, ninput), dtype=np.uint8)
    ngood = 0
    for pkt in packets:
        seq, _, npol, _, pn, _, _, chan0, pol0 = struct.unpack('>QLHHHHLLL', pkt[:32])
        if not (seq0 <= seq < seq0 + ntime):
            continue
        if npol <= 0 or pn <= 0 or chan0 < first_chan or chan0 + pn > first_chan + nchan or pol0 + npol > ninput:
            continue
        if 32 + pn * npol > len(pkt):
            continue
        data = np.frombuffer(pkt, dtype=np.uint8, count=pn * npol, offset=32).reshape(pn, npol)
        out[seq - seq0, chan0 - first_chan:chan0 - first_chan + pn, pol0:pol0 + npol] = data
        ngood += 1
    return out, ngood


def subsel_packet_payloads(sub, layout="py"):
    """CorrOutputPart's payloads from CorrSubsel's [nchan, nvis, 2] span.
    "py" : big-endian [nvis, nchan, 2] -- idata.transpose([1,0,2]) copied into dout (dtype '>i')
           (corr_output_part_block.py:429,460-463); returned as the raw bytes viewed as uint8.
    "cor": [nvis/4, nchan, 4, 2] -- dview = data.reshape([nchan, nvis//4, 2, 2]) then transpose([1,0,2,3])
           (corr_output_part_block.py:378-391)."""
    nchan, nvis, _ = sub.shape
    if layout == "py":
        dout = np.zeros([nvis, nchan, 2], dtype=">i")
        dout[...] = sub.transpose([1, 0, 2])
        return np.frombuffer(dout.tobytes(), dtype=np.uint8)
    dview = sub.reshape([nchan, nvis // 4, 2, 2, 2])
    return np.ascontiguousarray(dview.transpose([1, 0, 2, 3, 4])).reshape(nvis // 4, nchan, 4, 2)


def packet_payloads(reordered, layout="cor"):
    """Per-baseline payloads CorrOutputFull sends, from reorder()'s [s0, s1, p0, p1, chan, 2] matrix:
    baselines s0 <= s1 in the nested-loop order of send_packets_bf / send_packets_py.
    "cor": [nbl, nchan, npol, npol, 2] -- sdata.transpose([0,3,1,2,4]) (corr_output_full_block.py:505-511);
    "py" : [nbl, npol, npol, nchan, 2] -- reordered_data[s0, s1].tobytes() (corr_output_full_block.py:457-460)."""
    nstand = reordered.shape[0]
    rows = []
    for s0 in range(nstand):
        sdata = reordered[s0, s0:]                      # stand x pol x pol x chan x complexity
        rows.append(sdata.transpose([0, 3, 1, 2, 4]) if layout == "cor" else sdata)
    return np.ascontiguousarray(np.concatenate(rows, axis=0))


# ---------------------------------------------------------------------------
# Beamformer
# ---------------------------------------------------------------------------

def beamform(v_u8, weights):
    """out[c,b,t] = sum_i w[c,b,i] * v[t,c,i]  (no conjugation) --
    pipeline/verification/tests/beamformer/beamformer_test.py:76-84.
    v_u8: uint8 [ntime, nchan, ninput]; weights complex [nchan, nbeam, ninput].
    Returns complex128 [nchan, nbeam, ntime] (float64 arithmetic)."""
    r, i = unpack4(v_u8)
    v = r.astype(np.float64) + 1j * i.astype(np.float64)
    w = np.asarray(weights).astype(np.complex128)
    return np.einsum("cbi,tci->cbt", w, v, optimize=True)


def beam_weights(freqs, delays_ns, amps, cal_gains):
    """Beamform.update_command_vals, 'beamcoeffs' branch
    (beamform_block.py:343-348): weights for one beam, complex64
    [nchan, ninput]."""
    phases = np.exp(1j * 2 * np.pi * np.asarray(freqs)[:, None] * np.asarray(delays_ns) * 1e-9)
    return (np.asarray(amps) * phases * cal_gains).astype(np.complex64)


def beam_power_sum(x, ntime_sum):
    """bfBeamformIntegrate: pipeline/verification/tests/beamformer/
    beamformer_sum_test.py:64-77.  x: complex [nchan, nbeam, ntime].
    Returns float64 [nbeam//2, ntime//ntime_sum, nchan, 4] =
    [sum|X|^2, sum|Y|^2, Re sum X Y*, Im sum X Y*], X = beam 2b, Y = 2b+1."""
    nchan, nbeam, ntime = x.shape
    nb = ntime // ntime_sum
    d = np.asarray(x).astype(np.complex128)[:, :, : nb * ntime_sum]
    d = d.reshape(nchan, nbeam // 2, 2, nb, ntime_sum)
    X = d[:, :, 0]
    Y = d[:, :, 1]
    out = np.empty([nbeam // 2, nb, nchan, 4], dtype=np.float64)
    out[..., 0] = (np.abs(X) ** 2).sum(-1).transpose(1, 2, 0)
    out[..., 1] = (np.abs(Y) ** 2).sum(-1).transpose(1, 2, 0)
    cross = (X * np.conj(Y)).sum(-1).transpose(1, 2, 0)
    out[..., 2] = cross.real
    out[..., 3] = cross.imag
    return out


def beam_power_sum_single(x, ntime_sum, beam_id):
    """bfBeamformIntegrateSingleBeam (beamform_sum_single_beam_block.py:114):
    the dual-pol beam ``beam_id`` only -> [ntime//ntime_sum, nchan, 4]."""
    return beam_power_sum(x[:, 2 * beam_id : 2 * beam_id + 2, :], ntime_sum)[0]


# ---------------------------------------------------------------------------
# Offline upchannelise path
# ---------------------------------------------------------------------------

def upchannelize(v_u8, layout="fine_last"):
    """pipeline/scripts/lwa352-upchan-bf.py:94-96: TrigBufSourceBlock's ci4 [time][fine_time][freq][stand][pol]
    (trigger_source_block.py:37-41) is transposed to fine_time-last and Fourier transformed along fine_time by
    bifrost's fft block (forward, unnormalised, no fftshift -- stock bifrost, absent here; numpy's fft has the same
    definition).  v_u8: uint8 [nframe, nfine, nchan, ninput].  Returns complex128 [nframe, nchan, ninput, nfine]
    ("fine_last") or, as lwa352-upchan-imag.py:93-98 transposes it, [nframe, nchan, nfine, ninput]
    ("freq_merged")."""
    r, i = unpack4(v_u8)
    x = (r.astype(np.float64) + 1j * i.astype(np.float64)).transpose(0, 2, 3, 1)
    X = np.fft.fft(x, axis=-1)
    if layout == "freq_merged":
        X = X.transpose(0, 1, 3, 2)
    return np.ascontiguousarray(X)


def upchan_beamform(v_u8, weights):
    """lwa352-upchan-bf.py:94-113: upchannelise, multiply by cal * coeffs
    (beamform_offline_block.py:218-240; both [beam][chan][stand][pol][fine], constant over a gulp), reduce over
    stand then pol.  weights complex [nbeam, nchan, ninput, nfine].
    Returns complex128 [nframe, nbeam, nchan, nfine]."""
    X = upchannelize(v_u8)
    return np.einsum("bcik,fcik->fbck", np.asarray(weights).astype(np.complex128), X, optimize=True)


def upchan_correlate(X):
    """pipeline/scripts/lwa352-upchan-imag.py:93-106: bifrost's stock blocks.correlate on the upchannelised,
    frequency-merged voltages (absent submodule; conjugation as the reference's own correlator checks,
    check_golden_inputs.py:64-92 / corr_output_full_block.py:582-591: v_i * conj(v_j)).  X complex
    [nframe, nfreq, ninput].  Returns complex128 [nfreq, ninput, ninput] (the full Hermitian matrix)."""
    X = np.asarray(X).astype(np.complex128)
    return np.einsum("tfi,tfj->fij", X, np.conj(X), optimize=True)


def offline_beam_weights(freqs, delays_ns, weighting=None):
    """BfOfflineWeightsBlock.compute_weights, the part after the astropy pointing
    (beamform_offline_block.py:126-134): delays are made relative to the largest one, scaled by the per-input
    antenna weighting, and turned into phases at every fine frequency.  freqs [nchan * nfine] in Hz, delays_ns
    [ninput].  Returns complex [nchan * nfine, ninput]."""
    d = np.asarray(delays_ns, dtype=np.float64)
    d = d.max() - d
    if weighting is not None:
        d = np.asarray(weighting) * d
    return np.exp(2j * np.pi * np.asarray(freqs, dtype=np.float64)[:, None] * d * 1e-9)
