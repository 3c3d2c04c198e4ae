"""Host logic of the X-engine blocks (no GPU): integration state machines, ring headers, command handling.

The scenarios follow the behaviour documented in the reference blocks (corr_block.py:60-219,
corr_acc_block.py:17-168, corr_subsel_block.py:17-183, beamform_block.py:18-206) and the way its test scripts
drive them (verification/tests/beamformer/beamformer_test.py:176-182: JSON commands through
``process_command_strings``).  libxeng is replaced by tag-moving doubles (tests/fakes.py).
"""
import json
import logging

import numpy as np
import pytest

from xengine import ringshim
from xengine.blocks import (Beamform, BeamformSum, BeamformSumBeams, BeamformSumSingleBeam, Block, Corr, CorrAcc,
                            CorrSubsel)
from xengine.blocks import block_base

LOG = logging.getLogger("test")

NSTAND, NPOL, NCHAN, NT = 8, 2, 4, 4


def base_header(seq0=0, nchan=NCHAN, nstand=NSTAND, npol=NPOL, **extra):
    h = {'seq0': seq0, 'nchan': nchan, 'nstand': nstand, 'npol': npol, 'sfreq': 1000.0, 'bw_hz': 100.0 * nchan,
         'chan0': 0, 'fs_hz': 196000000, 'system_nchan': 32 * nchan, 'sync_time': 0, 'pipeline_id': 0,
         'ant_to_input': np.arange(nstand * npol).reshape(nstand, npol).tolist(),
         'input_to_ant': [[i // npol, i % npol] for i in range(nstand * npol)]}
    h.update(extra)
    return h


def fill(ring, sequences, time_tag0=0):
    """sequences: [(header dict, [span byte arrays])] -> written and closed."""
    with ring.begin_writing() as w:
        for k, (hdr, spans) in enumerate(sequences):
            with w.begin_sequence(time_tag=time_tag0 + k, header=json.dumps(hdr)) as seq:
                for s in spans:
                    s = np.ascontiguousarray(s).reshape(-1).view(np.uint8)
                    with seq.reserve(s.size) as span:
                        span.data[...] = s


def drain(ring, gulp):
    out = []
    for seq in ring.read(guarantee=False):
        hdr = json.loads(seq.header.tostring())
        spans = [np.array(fakes_np(sp.data)[:sp.size]) for sp in seq.read(gulp)]
        out.append((hdr, spans))
    return out


def fakes_np(buf):
    if hasattr(buf, "numpy"):
        buf = buf.numpy()
    return buf.reshape(-1).view(np.uint8)


def gulp(tag, nbytes):
    g = np.zeros(nbytes, dtype=np.uint8)
    g[0] = tag
    return g


# ------------------------------------------------------------------------------------------- Block base
class _Etcd(object):
    def __init__(self):
        self.puts, self.watch = [], None

    def add_watch_prefix_callback(self, key, cb):
        self.watch = (key, cb)
        return 7

    def put(self, key, val):
        self.puts.append((key, json.loads(val)))

    def cancel_watch(self, wid):
        self.cancelled = wid


def cmd(kwargs, seq_id="1", cmd="update"):
    return json.dumps({"id": seq_id, "cmd": cmd, "val": {"kwargs": kwargs}})


def test_block_command_validation_and_responses():
    etcd = _Etcd()
    Block.set_id(3)
    b = Block(LOG, ringshim.Ring("in"), ringshim.Ring("out"), True, -1, etcd_client=etcd, name="Thing")
    Block.set_id(0)
    assert b.command_key.startswith("/cmd/corr/x/") and b.command_key.endswith("/pipeline/3/Thing/%d" % b.instance_id)
    assert b.response_key.startswith("/resp/corr/x/") and b.monitor_key.startswith("/mon/corr/x/")
    assert etcd.watch[0] == b.command_key
    b.define_command_key("n", type=int, condition=lambda x: x % 2 == 0, initial_val=4)
    assert b.command_vals["n"] == 4 and not b.update_pending
    assert b._process_commands({"n": 6}) == block_base.COMMAND_OK and b.update_pending
    assert b.command_vals["n"] == 4                       # staged, not yet active
    b.update_command_vals()
    assert b.command_vals["n"] == 6 and not b.update_pending and b.stats["n"] == 6
    assert b._process_commands({"zzz": 1}) == block_base.COMMAND_NOT_RECOGNIZED
    assert b._process_commands({"n": "6"}) == block_base.COMMAND_WRONG_TYPE
    assert b._process_commands({"n": 5}) == block_base.COMMAND_INVALID
    assert not b.update_pending
    # the JSON envelope (block_base.py:216-255) and the response written to the /resp key
    b.process_command_strings(cmd({"n": 8}, seq_id="42"))
    key, resp = etcd.puts[-1]
    assert key == b.response_key and resp["id"] == "42" and resp["val"]["status"] == "normal"
    assert resp["val"]["response"] == "0" and b.update_pending
    b.process_command_strings([cmd({"n": 9}, seq_id="43"), json.dumps({"cmd": "update"}),
                               cmd({"n": 2}, cmd="bogus"), json.dumps({"id": "5", "cmd": "update", "val": 3}),
                               json.dumps({"id": "6", "cmd": "update", "val": {"kwargs": 1}})])
    got = [(r["id"], r["val"]["status"], r["val"]["response"]) for _, r in etcd.puts[-5:]]
    assert got == [("43", "error", "-3"), ("0", "error", "Missing ID field"), ("0", "error", "Invalid command"),
                   ("5", "error", "`val` field should be a dictionary"),
                   ("6", "error", "`val[kwargs]` field should be a dictionary")]
    assert b.stats["last_cmd_response"] == block_base.COMMAND_INVALID
    with pytest.raises(AssertionError):
        b.define_command_key("bad", type=int, initial_val="x")


def test_instance_ids_count_per_class(host_only):
    a = CorrAcc(LOG, ringshim.Ring(), ringshim.Ring(), nchan=NCHAN, nstand=NSTAND)
    b = CorrAcc(LOG, ringshim.Ring(), ringshim.Ring(), nchan=NCHAN, nstand=NSTAND)
    c = CorrSubsel(LOG, ringshim.Ring(), ringshim.Ring(), nchan=NCHAN, nstand=NSTAND)
    assert b.instance_id == a.instance_id + 1
    assert type(c)._instance_count == c.instance_id and type(a)._instance_count == b.instance_id
    assert a.name == "CorrAcc" and c.name == "CorrSubsel"


# ------------------------------------------------------------------------------------------- Corr
def make_corr(iring, oring, **kw):
    args = dict(ntime_gulp=NT, nchan=NCHAN, nstand=NSTAND, npol=NPOL, acc_len=2 * NT, autostartat=0)
    args.update(kw)
    return Corr(LOG, iring, oring, **args)


def test_corr_sizes_and_maps(host_only):
    c = make_corr(ringshim.Ring(), ringshim.Ring(),
                  ant_to_input=np.arange(NSTAND * NPOL).reshape(NSTAND, NPOL))
    wpc = (NSTAND // 2 + 1) * (NSTAND // 4) * 16
    assert c.matlen == NCHAN * wpc and c.igulp_size == NT * NCHAN * NSTAND * NPOL and c.ogulp_size == c.matlen * 8
    assert c.antpol_to_bl.shape == (NSTAND, NSTAND, NPOL, NPOL) and c.antpol_to_bl.dtype == np.int32
    # xeng_get_order is host-only index bookkeeping and runs without a GPU: check it against regtile_index
    from xengine.blocks.corr_block import regtile_index
    for s0 in range(NSTAND):
        for s1 in range(NSTAND):
            for p0 in range(2):
                for p1 in range(2):
                    if s1 >= s0:
                        assert c.antpol_to_bl[s0, s1, p0, p1] == regtile_index(2 * s0 + p0, 2 * s1 + p1, NSTAND)
                        assert c.bl_is_conj[s0, s1, p0, p1] == 1
                    else:
                        assert c.antpol_to_bl[s0, s1, p0, p1] == regtile_index(2 * s1 + p1, 2 * s0 + p0, NSTAND)
                        assert c.bl_is_conj[s0, s1, p0, p1] == 0
    with pytest.raises(AssertionError):
        make_corr(ringshim.Ring(), ringshim.Ring(), acc_len=NT + 1)


def test_corr_integrations_and_headers(host_only):
    iring, oring = ringshim.Ring("gpu-input"), ringshim.Ring("corr-output")
    c = make_corr(iring, oring, acc_len=2 * NT, autostartat=2 * NT)
    g = c.igulp_size
    fill(iring, [(base_header(0), [gulp(t, g) for t in range(1, 8)] + [np.zeros(g // 2, np.uint8)])])
    c.main()
    (hdr, spans), = drain(oring, c.ogulp_size)
    assert hdr['acc_len'] == 2 * NT and hdr['seq0'] == 2 * NT
    assert 'ant_to_input' not in hdr and 'input_to_ant' not in hdr and hdr['nchan'] == NCHAN
    # gulps 1,2 (samples 0..7) are before the start; 3+4 and 5+6 make two dumps; 7 is an unfinished integration
    tags = [tuple(s.view(np.int32)[:2]) for s in spans]
    assert tags == [(3 + 4, 2), (5 + 6, 2)]
    assert [x for x in host_only.CALLS if x[0] == "corr"] == [("corr", 3, True), ("corr", 4, False),
                                                              ("corr", 5, True), ("corr", 6, False),
                                                              ("corr", 7, True)]
    assert c.stats['state'] == 'running' and c.stats['last_end_sample'] == 5 * NT


def test_corr_start_asap_and_acc_len_command(host_only):
    iring, oring = ringshim.Ring(), ringshim.Ring()
    c = make_corr(iring, oring, acc_len=3 * NT, autostartat=-1)
    g = c.igulp_size
    # seq0 = 4*NT: next multiple of acc_len (12*NT/... ) after sample 16 is 24 -> gulps starting at 24, 28, 32
    fill(iring, [(base_header(4 * NT), [gulp(t, g) for t in range(1, 9)])])
    c.main()
    (hdr, spans), = drain(oring, c.ogulp_size)
    assert hdr['seq0'] == 6 * NT and hdr['acc_len'] == 3 * NT
    assert [tuple(s.view(np.int32)[:2]) for s in spans] == [(3 + 4 + 5, 3), (6 + 7 + 8, 3)]
    # command validation: multiples of the gulp only, -1 allowed for start_time
    assert c._process_commands({'acc_len': NT + 1}) == block_base.COMMAND_INVALID
    assert c._process_commands({'start_time': 3}) == block_base.COMMAND_INVALID
    assert c._process_commands({'start_time': -1, 'acc_len': 4 * NT}) == block_base.COMMAND_OK


def test_corr_realigns_on_new_sequence_and_stops_on_zero(host_only):
    iring, oring = ringshim.Ring(), ringshim.Ring()
    c = make_corr(iring, oring, acc_len=NT, autostartat=0)
    g = c.igulp_size
    # sequence 2 starts at sample 5*NT while running since 0: 5 integrations missed -> restart at (5+10)*NT
    fill(iring, [(base_header(0), [gulp(1, g), gulp(2, g)]),
                 (base_header(5 * NT), [gulp(10 + k, g) for k in range(12)])])
    c.main()
    out = drain(oring, c.ogulp_size)
    assert [h['seq0'] for h, _ in out] == [0, 15 * NT]
    assert [int(s.view(np.int32)[0]) for s in out[0][1]] == [1, 2]
    assert [int(s.view(np.int32)[0]) for s in out[1][1]] == [20, 21]      # gulps at samples 15*NT and 16*NT
    # acc_len = 0: stop, nothing is written, no crash
    iring, oring = ringshim.Ring(), ringshim.Ring()
    c = make_corr(iring, oring, acc_len=0, autostartat=0)
    fill(iring, [(base_header(0), [gulp(1, g), gulp(2, g)])])
    c.main()
    assert drain(oring, c.ogulp_size) == [] and c.stats['state'] == 'stopped'


def test_corr_fused_corracc_and_fast_dump_on_a_host_ring(host_only):
    """Corr(oring=None, acc_oring=..., slow_acc_len=..., fast_oring=<cuda_host ring>): one launch per gulp feeds the
    long accumulator (CorrAcc's span + header keys, corr_acc_block.py:216,255-256) and the fast dump (CorrSubsel's
    header keys, corr_subsel_block.py:272-285); the fast dump of a non-device ring goes through the device scratch."""
    iring, slow, fastr = ringshim.Ring("gpu-input"), ringshim.Ring("corr-slow-output"), ringshim.Ring("fast", space="cuda_host")
    c = make_corr(iring, None, acc_len=2 * NT, autostartat=0, acc_oring=slow, slow_acc_len=4 * NT, fast_oring=fastr,
                  nchan_sum=2, fast_nvis=6, ant_to_input=np.arange(NSTAND * NPOL).reshape(NSTAND, NPOL))
    assert c._fast_dev is not None
    g = c.igulp_size
    fill(iring, [(base_header(0), [gulp(t, g) for t in range(1, 11)])])
    c.main()
    (hdr, spans), = drain(slow, c.ogulp_size)
    assert hdr['upstream_acc_len'] == 2 * NT and hdr['acc_len'] == 4 * NT and hdr['seq0'] == 0
    assert [tuple(s.view(np.int32)[:2]) for s in spans] == [(1 + 2 + 3 + 4, 4), (5 + 6 + 7 + 8, 4)]   # 9, 10: unfinished
    (fhdr, fspans), = drain(fastr, c.fast_gulp_size)
    assert fhdr['nvis'] == 6 and fhdr['nchan'] == NCHAN // 2 and fhdr['nchan_sum'] == 2 and fhdr['acc_len'] == 2 * NT
    assert len(fhdr['baselines']) == 6
    assert [int(s.view(np.int32)[0]) for s in fspans] == [1 + 2, 3 + 4, 5 + 6, 7 + 8, 9 + 10]
    subs = [x for x in host_only.CALLS if x[0] == "corr_sub"]
    assert [x[2] for x in subs] == [True, False, False, False, True, False, False, False, True, False]    # slow first
    assert [x[3] for x in subs] == [True, False] * 5                                                      # fast first
    with pytest.raises(ValueError):
        make_corr(iring, ringshim.Ring(), acc_oring=slow, slow_acc_len=4 * NT)       # both outputs: not one launch
    with pytest.raises(ValueError):
        make_corr(iring, None, acc_len=2 * NT, acc_oring=slow, slow_acc_len=3 * NT)   # not a multiple of acc_len


def test_corr_multi_gulp_spans_sync_once_per_span(host_only):
    """gulps_per_span = 3: same dumps as one gulp per span, the device is waited for once per input span -- not once
    per gulp -- and the output spans of integrations that ended inside the span are committed after that wait; a short
    final span still yields its whole gulps."""
    outs = []
    for gps in (1, 3):
        del host_only.CALLS[:]
        iring, oring = ringshim.Ring(), ringshim.Ring()
        c = make_corr(iring, oring, acc_len=4 * NT, autostartat=0, gulps_per_span=gps)
        g = c.igulp_size
        fill(iring, [(base_header(0), [gulp(t, g) for t in range(1, 9)])])
        c.main()
        (hdr, spans), = drain(oring, c.ogulp_size)
        outs.append(([tuple(s.view(np.int32)[:2]) for s in spans],
                     [x[1:] for x in host_only.CALLS if x[0] == "corr"],
                     len([x for x in host_only.CALLS if x[0] == "corr_sync"])))
    assert outs[0][0] == outs[1][0] == [(1 + 2 + 3 + 4, 4), (5 + 6 + 7 + 8, 4)]
    assert outs[0][1] == outs[1][1]
    assert outs[0][2] == 8           # one wait per gulp
    assert outs[1][2] == 3           # input spans end after gulps 3, 6 and 8


# ------------------------------------------------------------------------------------------- DummySource / Copy
def test_dummy_source_header_ramp_and_copy_block(host_only, tmp_path):
    """DummySource -> Copy -> ring: the reference's fixture (dummy_source_block.py:194-195 ramp byte = stand % 8,
    :231-245 header) moved gulp by gulp by the Copy block (copy_block.py:92-160): headers unchanged, bytes
    unchanged, a partial final gulp ignored, buffer_multiplier gulps per input span."""
    from xengine.blocks import Copy, DummySource
    r0, r1 = ringshim.Ring("capture", space="system"), ringshim.Ring("gpu-input", space="system")
    src = DummySource(LOG, r0, ntime_gulp=NT, nchan=NCHAN, nstand=NSTAND, npol=NPOL, target_throughput=0,
                      pipeline_id=3, header={'extra': 1}, ngulp=5)
    assert src.gulp_size == NT * NCHAN * NSTAND * NPOL
    assert src.get_test_data()[0, 0, 5, 1] == 0 + 5j                  # byte 5 = re 0, im 5
    src.main()
    cp = Copy(LOG, r0, r1, ntime_gulp=NT, nbyte_per_time=NCHAN * NSTAND * NPOL, buffer_multiplier=2)
    assert cp.buf_size == 4 * 2 * src.gulp_size
    cp.main()
    (hdr, spans), = drain(r1, src.gulp_size)
    assert hdr['nchan'] == NCHAN and hdr['system_nchan'] == 32 * NCHAN and hdr['fs_hz'] == 196608000
    assert hdr['bw_hz'] == 24e3 * NCHAN and hdr['seq0'] == 0 and hdr['pipeline_id'] == 3 and hdr['extra'] == 1
    assert hdr['ant_to_input'] == np.arange(NSTAND * NPOL).reshape(NSTAND, NPOL).tolist()
    assert hdr['input_to_ant'][5] == [2, 1]
    assert len(spans) == 5
    want = np.broadcast_to((np.arange(NSTAND, dtype=np.uint8) % 8)[None, None, :, None], (NT, NCHAN, NSTAND, NPOL))
    for sp in spans:
        assert np.array_equal(sp.reshape(NT, NCHAN, NSTAND, NPOL), want)
    # test file source: gulps taken round-robin from the file; a short last input span is dropped by Copy
    raw = np.random.default_rng(0).integers(0, 256, size=2 * src.gulp_size, dtype=np.uint8)
    f = tmp_path / "vec.bin"
    raw.tofile(str(f))
    r0, r1 = ringshim.Ring(space="system"), ringshim.Ring(space="system")
    DummySource(LOG, r0, ntime_gulp=NT, nchan=NCHAN, nstand=NSTAND, npol=NPOL, target_throughput=0, testfile=str(f),
                ngulp=3).main()
    (_, got), = drain(r0, src.gulp_size)
    assert [bytes(g) for g in got] == [raw[:src.gulp_size].tobytes(), raw[src.gulp_size:].tobytes(), raw[:src.gulp_size].tobytes()]
    r2 = ringshim.Ring(space="system")
    fill(r2, [(base_header(0), [raw[:src.gulp_size], raw[:src.gulp_size // 2]])])
    Copy(LOG, r2, r1, ntime_gulp=NT, nbyte_per_time=NCHAN * NSTAND * NPOL).main()
    (_, got), = drain(r1, src.gulp_size)
    assert len(got) == 1 and bytes(got[0]) == raw[:src.gulp_size].tobytes()


# ------------------------------------------------------------------------------------------- CorrAcc
def corr_out_header(seq0, acc_len, **kw):
    h = base_header(seq0, acc_len=acc_len, **kw)
    h.pop('ant_to_input')
    h.pop('input_to_ant')
    return h


def vis_span(tag, nbytes):
    s = np.zeros(nbytes // 4, dtype=np.int32)
    s[0], s[1] = tag, 1
    return s


def test_corracc_accumulates_upstream_integrations(host_only):
    iring, oring = ringshim.Ring("corr-output"), ringshim.Ring("corr-slow-output", space="cuda_host")
    up = 2 * NT
    a = CorrAcc(LOG, iring, oring, nchan=NCHAN, nstand=NSTAND, acc_len=3 * up, autostartat=up)
    n = a.igulp_size
    assert a.ogulp_size == n == a.matlen * 8
    fill(iring, [(corr_out_header(0, up), [vis_span(t, n) for t in range(1, 9)])])
    a.main()
    (hdr, spans), = drain(oring, n)
    assert hdr['upstream_acc_len'] == up and hdr['acc_len'] == 3 * up and hdr['seq0'] == up
    # span 1 (sample 0) is before the start; 2+3+4 and 5+6+7 dump; 8 is unfinished
    assert [tuple(s.view(np.int32)[:2]) for s in spans] == [(9, 3), (18, 3)]
    acc_calls = [x for x in host_only.CALLS if x[0] == "acc"]
    assert acc_calls[:4] == [("acc", 2, True), ("acc", 3, False), ("acc", 4, False), ("acc", 5, True)]
    assert a.stats['last_end_sample'] == 7 * up - up


def test_corracc_start_now_realign_and_stop(host_only):
    up = NT
    iring, oring = ringshim.Ring(), ringshim.Ring()
    a = CorrAcc(LOG, iring, oring, nchan=NCHAN, nstand=NSTAND, acc_len=2 * up, autostartat=-1)
    n = a.igulp_size
    # -1 = start with the span in hand (corr_acc_block.py:245-246); a new sequence at 7*up realigns to
    # start + (missed + 2) * acc_len = 3*up + (2 + 2) * 2*up = 11*up
    fill(iring, [(corr_out_header(3 * up, up), [vis_span(t, n) for t in (1, 2, 3)]),
                 (corr_out_header(7 * up, up), [vis_span(10 + k, n) for k in range(7)])])
    a.main()
    out = drain(oring, n)
    assert [h['seq0'] for h, _ in out] == [3 * up, 11 * up]
    assert [int(s.view(np.int32)[0]) for s in out[0][1]] == [3]
    assert [int(s.view(np.int32)[0]) for s in out[1][1]] == [14 + 15]
    iring, oring = ringshim.Ring(), ringshim.Ring()
    a = CorrAcc(LOG, iring, oring, nchan=NCHAN, nstand=NSTAND, acc_len=0, autostartat=0)
    fill(iring, [(corr_out_header(0, up), [vis_span(1, n)])])
    a.main()
    assert drain(oring, n) == [] and a.stats['state'] == 'stopped'


# ------------------------------------------------------------------------------------------- CorrSubsel
def test_corrsubsel_headers_selection_and_mid_sequence_update(host_only, monkeypatch):
    monkeypatch.setattr(CorrSubsel, "nvis_out", 6)
    a2bl = np.arange(NSTAND * NSTAND * 4, dtype=np.int32).reshape(NSTAND, NSTAND, 2, 2)
    conj = (a2bl % 3 == 0).astype(np.int32)
    iring, oring = ringshim.Ring("corr-output"), ringshim.Ring("corr-fast-output", space="cuda_host")
    s = CorrSubsel(LOG, iring, oring, nchan=NCHAN, nstand=NSTAND, nchan_sum=2, antpol_to_bl=a2bl, bl_is_conj=conj)
    assert s.nchan_out == NCHAN // 2 and s.ogulp_size == s.nchan_out * 6 * 8
    # default selection: pol-0 autos of stand v % nstand
    assert list(fakes_np(s._subsel_next).view(np.int32)) == [int(a2bl[v, v, 0, 0]) for v in range(6)]
    bls = [[[1, 0], [2, 1]], [[2, 1], [1, 0]], [[0, 0], [0, 0]], [[7, 1], [3, 0]], [[3, 0], [7, 1]], [[5, 1], [5, 0]]]
    assert s._process_commands({'baselines': bls[:3]}) == block_base.COMMAND_INVALID      # wrong length
    assert s._process_commands({'baselines': bls}) == block_base.COMMAND_OK
    n = s.igulp_size
    fill(iring, [(corr_out_header(100, 10), [vis_span(t, n) for t in (1, 2, 3)])])
    # a second selection arrives while the sequence is being processed: injected after the first span
    newbls = [[[v, 1], [v, 1]] for v in range(6)]
    real = host_only.fake_subsel
    state = {'n': 0}

    def subsel_then_command(*a, **k):
        real(*a, **k)
        state['n'] += 1
        if state['n'] == 1:
            s.process_command_strings(cmd({'baselines': newbls}))

    from xengine import ops
    monkeypatch.setattr(ops, "subsel", subsel_then_command)
    s.main()
    out = drain(oring, s.ogulp_size)
    assert len(out) == 2
    h0, h1 = out[0][0], out[1][0]
    assert h0['nchan'] == NCHAN // 2 and h0['nvis'] == 6 and h0['nchan_sum'] == 2 and h0['baselines'] == bls
    chan_width = 100.0
    assert h0['sfreq'] == (1000.0 + (2 - 1) * chan_width) / 2          # corr_subsel_block.py:274-275
    assert h0['seq0'] == 100 and h1['seq0'] == 110 and h1['baselines'] == newbls
    # first span used the first selection, the next two the new one (table entry 0 identifies the table)
    first_tab = int(a2bl[1, 2, 0, 1])
    new_tab = int(a2bl[0, 0, 1, 1])
    assert [tuple(sp.view(np.int32)[:2]) for sp in out[0][1]] == [(1, first_tab)]
    assert [tuple(sp.view(np.int32)[:2]) for sp in out[1][1]] == [(2, new_tab), (3, new_tab)]
    assert list(fakes_np(s._conj).view(np.int32)) == [int(conj[v, v, 1, 1]) for v in range(6)]


# ------------------------------------------------------------------------------------------- Beamform
NBEAM, NINPUT = 4, NSTAND * NPOL


def calgain_cmd(beam, inp, gains, seq_id="1"):
    data = np.stack([gains.real, gains.imag], axis=1).reshape(-1).tolist()
    return cmd({'coeffs': {'type': 'calgains', 'input_id': inp, 'beam_id': beam, 'data': data}}, seq_id)


def beam_cmd(beam, delays, amps, load_sample=None, seq_id="2"):
    d = {'type': 'beamcoeffs', 'beam_id': beam, 'data': {'delays': list(delays), 'amps': list(amps)}}
    if load_sample is not None:
        d['load_sample'] = load_sample
    return cmd({'coeffs': d}, seq_id)


def test_corr_output_full_packets_headers_and_destination(host_only, tmp_path):
    """CorrOutputFull host logic (corr_output_full_block.py:440-548,604-698): one packet per baseline s0 <= s1 in
    nested-loop order, COR header fields / plain header fields, time tags advancing by acc_len, the tuning word,
    `dest_file` command opening a file that receives header + payload back to back."""
    import struct
    from xengine.blocks import CorrOutputFull
    matlen = NCHAN * (NSTAND // 2 + 1) * (NSTAND // 4) * NPOL * NPOL * 4
    nbl = NSTAND * (NSTAND + 1) // 2
    a2bl = np.arange(NSTAND * NSTAND * 4, dtype=np.int32).reshape(NSTAND, NSTAND, 2, 2) + 100
    for cor in (True, False):
        del host_only.CALLS[:]
        ring = ringshim.Ring("slow", space="system")
        hdr = base_header(seq0=40, acc_len=20, upstream_acc_len=10)
        fill(ring, [(hdr, [gulp(3, matlen * 8), gulp(5, matlen * 8), gulp(9, 16)])])     # third span is short: skipped
        blk = CorrOutputFull(LOG, ring, guarantee=False, nchan=NCHAN, npol=NPOL, nstand=NSTAND, antpol_to_bl=a2bl,
                             bl_is_conj=np.zeros_like(a2bl), use_cor_fmt=cor, nchan_sum=4, pipeline_idx=5, npipeline=4)
        assert blk.tuning == (4 << 16) | (4 << 8) | 1 and blk.igulp_size == matlen * 8
        assert blk.dump_size == NSTAND * (NSTAND + 1) * NPOL * NPOL * NCHAN * 2 * 4 // 2
        pkts = []
        blk.sink = lambda h, p: pkts.append((h, p.copy()))
        blk.main()
        assert [c[:3] for c in host_only.CALLS] == [("reorder_packets", 3, 1 if cor else 2), ("reorder_packets", 5, 1 if cor else 2)]
        assert host_only.CALLS[0][3] == 101                                   # the maps given to the ctor were uploaded
        assert len(pkts) == 2 * nbl
        assert pkts[0][1][0] == 3 and pkts[nbl][1][0] == 5                    # payload rows come from the right dump
        assert pkts[0][1].size == NCHAN * NPOL * NPOL * 2
        order = [(s0, s1) for s0 in range(NSTAND) for s1 in range(s0, NSTAND)]
        spectra = int(NCHAN * hdr['fs_hz'] / hdr['bw_hz'])
        for k, (h, _) in enumerate(pkts):
            dump, (s0, s1) = divmod(k, nbl)[0], order[k % nbl]
            t = 40 + 20 * dump
            if cor:
                sync, idf, tun, ch0, gain, tt, navg, si, sj = struct.unpack(">IIIhhqihh", h)
                assert (sync, idf >> 24, idf & 0xFFFFFF, tun) == (0x5CDEC0DE, 2, nbl, blk.tuning)
                assert (ch0, gain, tt, navg, si, sj) == (0, 0, t * spectra, 20 * spectra, s0 + 1, s1 + 1)
            else:
                f = struct.unpack(">QQ2d4I2I", h)
                assert f == (0, t, hdr['bw_hz'], hdr['sfreq'], 20, NCHAN, 0, NPOL, s0, s1)
    # destination file through the command interface
    ring = ringshim.Ring("slow", space="system")
    fill(ring, [(base_header(seq0=0, acc_len=20), [gulp(7, matlen * 8)])])
    blk = CorrOutputFull(LOG, ring, guarantee=False, nchan=NCHAN, npol=NPOL, nstand=NSTAND, antpol_to_bl=a2bl,
                         bl_is_conj=np.zeros_like(a2bl))
    path = tmp_path / "cor.bin"
    blk.process_command_strings(cmd({"dest_file": str(path)}))
    blk.main()
    raw = path.read_bytes()
    assert len(raw) == nbl * (32 + NCHAN * NPOL * NPOL * 8)
    assert struct.unpack(">hh", raw[28:32]) == (1, 1) and struct.unpack("<i", raw[32:36]) == (7,)


def test_beamform_weights_and_load_sample(host_only):
    rng = np.random.default_rng(0xaabbccdd)
    iring, oring = ringshim.Ring("gpu-input"), ringshim.Ring("bf-output")
    b = Beamform(LOG, iring, oring, nchan=NCHAN, nbeam=NBEAM, ninput=NINPUT, ntime_gulp=NT)
    g = NT * NCHAN * NINPUT
    hdr = base_header(0)
    freqs = hdr['sfreq'] + (hdr['bw_hz'] / NCHAN) * np.arange(NCHAN)
    cal = (3 * rng.random((NCHAN, NINPUT)) + 4) + 1j * (4 * rng.random((NCHAN, NINPUT)) + 5)
    delays = 12 * rng.random(NINPUT)
    amps = 7 * rng.random(NINPUT) + 10
    state = {'n': 0}
    real_run = host_only.FakeBeamformer.run

    def run_then_command(self, d_in, d_out, d_w):
        real_run(self, d_in, d_out, d_w)
        state['n'] += 1
        if state['n'] == 1:       # after gulp 0: calibrate beam 1, then point it (immediately) ...
            b.process_command_strings([calgain_cmd(1, i, cal[:, i]) for i in range(NINPUT)])
            b.process_command_strings(beam_cmd(1, delays, amps))
            # ... and schedule beam 2 for sample 3*NT
            b.process_command_strings(beam_cmd(2, delays * 0, amps * 0 + 1, load_sample=3 * NT))

    host_only.FakeBeamformer.run = run_then_command
    try:
        fill(iring, [(hdr, [gulp(t, g) for t in range(1, 6)])])
        b.main()
    finally:
        host_only.FakeBeamformer.run = real_run
    (ohdr, spans), = drain(oring, NT * NCHAN * NBEAM * 8)
    assert (ohdr['nstand'], ohdr['nbeam'], ohdr['nbit'], ohdr['npol'], ohdr['complex']) == (NBEAM, NBEAM, 32, 1, True)
    assert len(spans) == 5
    runs = [x for x in host_only.CALLS if x[0] == "bf_run"]
    w = [r[2].reshape(NCHAN, NBEAM, NINPUT) for r in runs]
    assert not w[0].any()                                           # gulp 0: no coefficients loaded yet
    want1 = (amps * np.exp(1j * 2 * np.pi * freqs[:, None] * delays * 1e-9) * cal).astype(np.complex64)
    for k in (1, 2, 3, 4):                                          # from gulp 1 on, beam 1 is loaded
        np.testing.assert_allclose(w[k][:, 1, :], want1, rtol=1e-6)
        assert not w[k][:, 0, :].any() and not w[k][:, 3, :].any()
    assert not w[2][:, 2, :].any()                                  # gulp 2 starts at 2*NT < load_sample
    np.testing.assert_allclose(w[3][:, 2, :], np.ones((NCHAN, NINPUT)), rtol=1e-6)   # gulp 3 starts at 3*NT
    assert b.stats['cal_gains1'] == [True] * NINPUT and b.stats['cal_gains0'] == [False] * NINPUT
    assert not b.update_pending


def test_beamform_experimental_power_mode(host_only):
    iring, oring = ringshim.Ring(), ringshim.Ring()
    b = Beamform(LOG, iring, oring, nchan=NCHAN, nbeam=NBEAM, ninput=NINPUT, ntime_gulp=NT, ntime_sum=2)
    assert b.ntime_blocks == NT // 2 and b._bf.dims[-1] == NT // 2
    g = NT * NCHAN * NINPUT
    fill(iring, [(base_header(0), [gulp(1, g), gulp(2, g)])])
    b.main()
    (ohdr, spans), = drain(oring, (NT // 2) * NCHAN * NBEAM * 8)
    assert len(spans) == 2 and [c[:3] for c in host_only.CALLS if c[0].startswith("bf_run")] == \
        [("bf_run_integrate", 1, 2), ("bf_run_integrate", 2, 2)]
    with pytest.raises(AssertionError):
        Beamform(LOG, iring, oring, nchan=NCHAN, nbeam=NBEAM, ninput=NINPUT, ntime_gulp=NT, ntime_sum=3)


# ------------------------------------------------------------------------------------------- BeamformSum*
def bf_header(nbeam, **kw):
    h = base_header(0, **kw)
    h.update({'nstand': nbeam, 'nbeam': nbeam, 'nbit': 32, 'npol': 1, 'complex': True})
    return h


def test_beamform_sum_blocks_headers_and_sizes(host_only):
    nt, ns = 8, 4
    ig = nt * NCHAN * NBEAM * 8
    iring, oring = ringshim.Ring("bf-output"), ringshim.Ring("bf-pow-output", space="cuda_host")
    b = BeamformSumBeams(LOG, iring, oring, nchan=NCHAN, ntime_gulp=nt, ntime_sum=ns)
    fill(iring, [(bf_header(NBEAM), [gulp(1, ig), gulp(2, ig), np.zeros(ig // 2, np.uint8)])])
    b.main()
    og = (nt // ns) * NCHAN * (NBEAM // 2) * 16
    (h, spans), = drain(oring, og)
    assert (h['nbeam'], h['nbit'], h['complex'], h['acc_len'], h['npol']) == (NBEAM // 2, 32, True, ns, 2)
    assert [len(s) for s in spans] == [og, og] and [int(s[0]) for s in spans] == [1, 2]
    assert [c for c in host_only.CALLS if c[0] == "bf_integrate"] == [("bf_integrate", 1, ns), ("bf_integrate", 2, ns)]

    iring, oring = ringshim.Ring(), ringshim.Ring()
    b = BeamformSumSingleBeam(LOG, iring, oring, nchan_max=NCHAN, ntime_gulp=nt, ntime_sum=ns, beam_id=1)
    fill(iring, [(bf_header(NBEAM), [gulp(5, ig)])])
    b.main()
    og1 = (nt // ns) * NCHAN * 16
    (h, spans), = drain(oring, og1)
    assert (h['nstand'], h['nbeam'], h['beam_id'], h['ntime_block'], h['npol'], h['acc_len']) == (1, 1, 1, nt // ns, 2, ns)
    assert [len(s) for s in spans] == [og1]
    assert host_only.CALLS[-1] == ("bf_integrate_single", 5, ns, 1)

    iring, oring = ringshim.Ring(), ringshim.Ring()
    b = BeamformSum(LOG, iring, oring, nchan_max=NCHAN, nbeam_max=NBEAM, nstand=NSTAND, ntime_gulp=nt, ntime_sum=ns)
    fill(iring, [(bf_header(NBEAM), [gulp(9, ig)])])
    b.main()
    og2 = (nt // ns) * NCHAN * NBEAM * 16                       # the reference's (over-allocated) span size
    (h, spans), = drain(oring, og2)
    assert (h['nstand'], h['nbeam'], h['npol'], h['ntime_block']) == (1, NBEAM // 2, 2, nt // ns)
    assert [len(s) for s in spans] == [og2]
    with pytest.raises(AssertionError):
        BeamformSumBeams(LOG, iring, oring, nchan=NCHAN, ntime_gulp=10, ntime_sum=4)


def test_beamform_output_packets_and_destinations(host_only):
    """BeamformOutput host logic (beamform_output_block.py:243-245,311-317,353-361): destinations assigned to beams
    round-robin from the command lists, one packet per (beam, time step) carrying the contiguous [nchan][4] row, the
    header bytes of send_packets_python, seq advancing by ntime_gulp * navg per packet, beams at 0.0.0.0 skipped."""
    import struct
    from xengine.blocks import BeamformOutput
    nt, nbeam = 3, 4
    pw = np.arange(nbeam * nt * NCHAN * 4, dtype=np.float32)
    ring = ringshim.Ring("bf-pow")
    hdr = bf_header(nbeam, acc_len=24)
    hdr.update({'npol': 2, 'chan0': 96, 'seq0': 480})
    fill(ring, [(hdr, [pw, pw[:8]])])                                     # second span is short: skipped
    blk = BeamformOutput(LOG, ring, guarantee=False, ntime_gulp=nt, pipeline_idx=3, nchan=NCHAN, nbeam=nbeam)
    pkts = []
    blk.sink = lambda b, h, p: pkts.append((b, h, p))
    blk.process_command_strings(cmd({"dest_ip": ["10.0.0.1", "0.0.0.0"], "dest_port": [10000, 10001, 10002]}))
    blk.main()
    assert blk.beam_ips == ["10.0.0.1", "0.0.0.0", "10.0.0.1", "0.0.0.0"] and blk.beam_ports == [10000, 10001, 10002, 10000]
    assert [p[0] for p in pkts] == [0] * nt + [2] * nt                    # beams 1 and 3 have no destination
    rows = pw.reshape(nbeam, nt, NCHAN * 4)
    for k, (b, h, p) in enumerate(pkts):
        t = k % nt
        assert struct.unpack(">6B2HQ", h) == (2, b, 1, NCHAN, nbeam, 32, 24, 96, 480 + t * nt * 24)
        assert np.array_equal(np.frombuffer(p, dtype=np.float32), rows[b, t])


def test_beamform_vlbi_output_rows(host_only):
    """BeamformVlbiOutput host logic (beamform_vlbi_output_block.py:228,260-262): span cf32 [nchan][nbeam][ntime];
    the first two single-pol beams go out as one [nchan][2] row per time sample, seq = seq0 + sample index; nothing
    is sent while dest_ip is 0.0.0.0."""
    from xengine.blocks import BeamformVlbiOutput
    nt, nbeam = 40, 4
    rng = np.random.default_rng(5)
    volt = (rng.standard_normal((NCHAN, nbeam, nt)) + 1j * rng.standard_normal((NCHAN, nbeam, nt))).astype(np.complex64)
    for ip, nexp in (("0.0.0.0", 0), ("10.0.0.9", 2 * nt)):
        ring = ringshim.Ring("bf-volt")
        vh = bf_header(nbeam)
        vh['seq0'] = 1000
        fill(ring, [(vh, [volt, volt * 2])])
        blk = BeamformVlbiOutput(LOG, ring, guarantee=False, ntime_gulp=nt, pipeline_idx=1)
        rows = []
        blk.sink = lambda seq, p: rows.append((seq, np.frombuffer(p, dtype=np.complex64)))
        blk.process_command_strings(cmd({"dest_ip": ip}))
        blk.main()
        assert len(rows) == nexp
        for k, (seq, r) in enumerate(rows):
            g, t = divmod(k, nt)
            assert seq == 1000 + k
            assert np.array_equal(r.reshape(NCHAN, 2), (volt * (g + 1))[:, 0:2, t])
