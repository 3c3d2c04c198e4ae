#!/usr/bin/env python
"""
Generate tests/golden/*.npz by EXECUTING THE REFERENCE'S OWN numpy code.

Runs only in the build container (needs /root/reference); the fixtures it
writes are committed, so nothing under tests/ touches /root/reference at run
time.  The reference modules import bifrost / ujson / etcd3 / simplejson, which
are absent here: those imports are satisfied by inert stub modules (a mock for
the runtime, the stdlib ``json`` for the JSON packages) so that the pure-numpy
functions we want can be called unmodified.  ``np.complex`` (removed in numpy
>= 1.24, used by the reference) is aliased to ``complex``.

What is executed from the reference:
  * pipeline/verification/make_golden_inputs.py (whole script, random+chanramp)
      -> golden_corr_random.npz, golden_corr_chanramp.npz
  * pipeline/lwa352_pipeline/blocks/corr_block.py: tri_index, regtile_index,
    Corr._test                      -> golden_regtile.npz, golden_corr_test.npz
  * pipeline/verification/tests/beamformer/beamformer_test.py:
    SoftwareBf.cpu_beamform / get_input_gulp, and the weight formula in main()
  * pipeline/lwa352_pipeline/blocks/beamform_block.py:
    Beamform.update_command_vals driven by
    pipeline-control/.../beamform_control.py JSON commands
                                    -> golden_beamform.npz
  * pipeline/verification/tests/beamformer/beamformer_sum_test.py:
    SoftwareBfSum.cpu_sum_power     -> golden_beamsum.npz
  * pipeline/test_transmitters/test_tx_vectors.py (whole script, its UDP socket replaced by a recorder)
                                    -> golden_snap2_packets.npz

Usage:  python tests/golden/make_golden.py
"""
import hashlib
import importlib
import importlib.abc
import importlib.machinery
import importlib.util
import json
import logging
import os
import runpy
import sys
import tempfile
import types
from unittest import mock

import numpy as np

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))


class _StubFinder(importlib.abc.MetaPathFinder, importlib.abc.Loader):
    """Satisfy ``import bifrost[.anything]`` / ``import etcd3`` with inert mocks."""
    ROOTS = ("bifrost", "etcd3")

    def find_spec(self, fullname, path, target=None):
        if fullname.split(".")[0] in self.ROOTS:
            return importlib.machinery.ModuleSpec(fullname, self, is_package=True)
        return None

    def create_module(self, spec):
        m = mock.MagicMock(name=spec.name)
        m.__name__ = spec.name
        m.__path__ = []
        m.__spec__ = spec
        m.__loader__ = self
        return m

    def exec_module(self, module):
        pass


def install_stubs():
    if not hasattr(np, "complex"):
        np.complex = complex  # reference predates numpy 1.24
    sys.meta_path.insert(0, _StubFinder())
    sys.modules["ujson"] = json
    sys.modules["simplejson"] = json
    sys.path.insert(0, os.path.join(REF, "pipeline"))
    sys.path.insert(0, os.path.join(REF, "pipeline-control"))


class FakeDeviceArray(np.ndarray):
    """Stands in for a bifrost ndarray: .copy(space=...) returns host data."""

    def copy(self, space=None, order="C"):
        return np.array(self, copy=True).view(FakeDeviceArray)


def load_ref_module(relpath, modname):
    """Import a reference *script* (not on a package path) as a module."""
    spec = importlib.util.spec_from_file_location(modname, os.path.join(REF, relpath))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def run_make_golden_inputs(extra):
    """Execute the reference fixture generator unmodified; return (in, corr)."""
    with tempfile.TemporaryDirectory() as td:
        argv = ["make_golden_inputs.py", "--datapath", td] + extra
        with mock.patch.object(sys, "argv", argv):
            try:
                runpy.run_path(os.path.join(REF, "pipeline/verification/make_golden_inputs.py"),
                               run_name="__main__")
            except SystemExit:
                pass
        fin = [f for f in os.listdir(td) if f.startswith("in_")][0]
        fco = [f for f in os.listdir(td) if f.startswith("corr_")][0]
        with open(os.path.join(td, fin), "rb") as fh:
            hin = json.loads(fh.readline().decode())
            din = np.frombuffer(fh.read(), dtype=np.uint8).reshape(hin["shape"])
        with open(os.path.join(td, fco), "rb") as fh:
            hco = json.loads(fh.readline().decode())
            dco = np.frombuffer(fh.read(), dtype=complex).reshape(hco["shape"])
    return hin, din, hco, dco


def main():
    install_stubs()
    log = logging.getLogger("golden")

    # ---- 1. reference fixture generator, random (seed 0xdeadbeef) + chanramp
    for tag, extra in [("random", []), ("chanramp", ["--chanramp"])]:
        hin, din, hco, dco = run_make_golden_inputs(
            ["-t", "64", "-c", "3", "-s", "12", "-p", "2", "--accshort", "32"] + extra)
        assert np.all(dco.real == np.rint(dco.real)) and np.all(dco.imag == np.rint(dco.imag))
        np.savez_compressed(os.path.join(HERE, "golden_corr_%s.npz" % tag),
                            vin=din, corr_re=dco.real.astype(np.int32),
                            corr_im=dco.imag.astype(np.int32), acc_len=hco["acc_len"],
                            seed=hin["seed"])
        print(tag, din.shape, dco.shape)

    # ---- 2. corr_block.py: index functions and the in-line numpy correlator
    cb = importlib.import_module("lwa352_pipeline.blocks.corr_block")
    out = {}
    for nstand in (4, 8, 12, 352):
        ni = 2 * nstand
        in0, in1 = np.meshgrid(np.arange(ni), np.arange(ni), indexing="ij")
        ok = (in1 >> 1) >= (in0 >> 1)
        idx = np.array([cb.regtile_index(int(a), int(b), nstand)
                        for a, b in zip(in0[ok], in1[ok])], dtype=np.int64)
        if nstand == 352:
            # 248512 entries: keep a digest plus a strided sample
            out["regtile_352_sha256"] = np.frombuffer(
                hashlib.sha256(idx.astype("<i8").tobytes()).digest(), dtype=np.uint8)
            out["regtile_352_sample_pos"] = np.arange(0, idx.size, 997)
            out["regtile_352_sample"] = idx[::997]
            out["regtile_352_max"] = idx.max()
        else:
            out["regtile_%d" % nstand] = idx
    out["tri"] = np.array([[cb.tri_index(i, j) for j in range(8)] for i in range(8)])
    np.savez_compressed(os.path.join(HERE, "golden_regtile.npz"), **out)

    rng = np.random.RandomState(0xdeadbeef)
    nt, nchan, nstand, npol = 16, 2, 8, 2
    vin = rng.randint(0, 256, [nt, nchan, nstand, npol]).astype(np.uint8)
    vin.flat[:6] = [0xFF, 0x88, 0x80, 0x08, 0x78, 0x87]  # edge bytes randint(0,255) never gives
    fake_self = types.SimpleNamespace(ntime_gulp=nt)
    ref_out = cb.Corr._test(fake_self, vin.view(FakeDeviceArray), nchan, nstand, npol)
    np.savez_compressed(os.path.join(HERE, "golden_corr_test.npz"), vin=vin,
                        out_re=ref_out.real.astype(np.int32), out_im=ref_out.imag.astype(np.int32))
    print("Corr._test", ref_out.shape)

    # ---- 3. beamformer: SoftwareBf + Beamform.update_command_vals
    bt = load_ref_module("pipeline/verification/tests/beamformer/beamformer_test.py", "ref_bf_test")
    from lwa352_pipeline.blocks.beamform_block import Beamform
    from lwa352_pipeline_control.blocks.beamform_control import BeamformControl
    from lwa352_pipeline_control.etcd_control import EtcdCorrControl

    NSTAND, NPOL, NCHAN, NT, NBEAM = 6, 2, 4, 10, 2   # 2*NBEAM single-pol beams
    ninput = NSTAND * NPOL
    rng = np.random.default_rng(0xaabbccdd)             # beamformer_test.py:121
    test_data = rng.integers(0, high=255, size=[NT, NCHAN, NSTAND, NPOL], dtype=np.uint8)
    # weight recipe of beamformer_test.py:131-139
    calgains = (3 * rng.random(size=[NCHAN, 2 * NBEAM, ninput]) + 4) \
        + 1j * (4 * rng.random(size=[NCHAN, 2 * NBEAM, ninput]) + 5)
    calgains = np.array(calgains, dtype=np.complex64)
    beamdelays = 12 * rng.random(size=[2 * NBEAM, ninput], dtype=np.float32)
    beamamps = 7 * rng.random(size=[2 * NBEAM, ninput], dtype=np.float32) + 10
    chan_bw_hz = bt.chan_bw_hz
    coeffs = np.zeros([NCHAN, 2 * NBEAM, ninput], dtype=np.complex64)
    for b in range(2 * NBEAM):
        for i in range(ninput):
            coeffs[:, b, i] = calgains[:, b, i] * beamamps[b, i] * np.exp(
                1j * 2 * np.pi * beamdelays[b, i] / 1e9 * chan_bw_hz * np.arange(NCHAN))

    with tempfile.NamedTemporaryFile(suffix=".dat") as tf:
        tf.write(test_data.tobytes())
        tf.flush()
        sbf = bt.SoftwareBf(log, iring=mock.MagicMock(), ntime_gulp=NT, testfile=tf.name, coeffs=coeffs)
        cdata = sbf.get_input_gulp(0)
        ref_bf = sbf.cpu_beamform(cdata)
        sbf.testfile.close()

    # The reference block turns control JSON into weights: run that code too.
    ctl_if = EtcdCorrControl(simulated=True, log=log)
    ctl = BeamformControl(log, ctl_if, "")
    blk = Beamform(log, iring=mock.MagicMock(), oring=mock.MagicMock(), nchan=NCHAN,
                   nbeam=2 * NBEAM, ninput=ninput, ntime_gulp=NT)
    blk.freqs = chan_bw_hz * np.arange(NCHAN)
    cmds = []
    for b in range(2 * NBEAM):
        for i in range(ninput):
            cmds.append(ctl.update_calibration_gains(b, i, calgains[:, b, i]))
    for b in range(2 * NBEAM):
        cmds.append(ctl.update_delays(b, beamdelays[b], beamamps[b]))
    blk.process_command_strings(cmds)
    np.savez_compressed(os.path.join(HERE, "golden_beamform.npz"), vin=test_data,
                        coeffs=coeffs, calgains=calgains, delays=beamdelays, amps=beamamps,
                        chan_bw_hz=chan_bw_hz, unpacked=np.asarray(cdata).astype(np.complex64),
                        out=np.asarray(ref_bf).astype(np.complex64),
                        block_gains=np.asarray(blk.gains_cpu_new).astype(np.complex64),
                        commands=np.array(cmds[:1] + cmds[-1:]))
    ExcEtcd = getattr(ctl_if, "ec", None)
    ctl_if.ec = mock.MagicMock()   # its __del__ closes the (absent) etcd client
    print("beamform", ref_bf.shape, "cmd example:", cmds[-1][:120])

    # ---- 4. beam power sum: SoftwareBfSum.cpu_sum_power
    bs = load_ref_module("pipeline/verification/tests/beamformer/beamformer_sum_test.py", "ref_bfsum_test")
    rng = np.random.default_rng(0xaabbccdd)             # beamformer_sum_test.py:116
    nchan, nbeam, ntime, ntime_sum = 3, 4, 48, 24
    x = (rng.random(size=[nchan, nbeam, ntime]) - 0.5) + 1j * (rng.random(size=[nchan, nbeam, ntime]) - 0.5)
    x = np.array(x, dtype=np.complex64) * 5
    sb = bs.SoftwareBfSum(log, iring=mock.MagicMock(), ntime_gulp=ntime)
    ref_pow = sb.cpu_sum_power(x, ntime_sum)
    np.savez_compressed(os.path.join(HERE, "golden_beamsum.npz"), x=x, ntime_sum=ntime_sum, out=ref_pow)
    print("beamsum", ref_pow.shape)


def snap2_packets():
    """Run the reference's SNAP2 emulator (pipeline/test_transmitters/test_tx_vectors.py) unmodified on a small vector
    file; its socket is replaced by a recorder that raises KeyboardInterrupt (which the script catches to stop) after
    every spectrum of the file has been sent once."""
    rng = np.random.RandomState(0x5a92)
    ntime, nchan, nstand, npol = 6, 8, 64, 2
    data = rng.randint(0, 256, [ntime, nchan, nstand, npol]).astype(np.uint8)
    hdr = {"dtype": "np.uint8", "shape": list(data.shape), "ntime": ntime, "nchan": nchan, "nstand": nstand,
           "npol": npol}
    npkt = ntime * 2 * (nstand // 32)                      # -b 2 channel blocks x (nstand / 32) stand blocks
    sent = []

    class Recorder(object):
        def __init__(self, *a, **k):
            pass

        def sendto(self, payload, addr):
            if len(sent) >= npkt:
                raise KeyboardInterrupt
            sent.append(bytes(payload))

    with tempfile.NamedTemporaryFile(suffix=".dat") as tf:
        tf.write((json.dumps(hdr) + "\n").encode())
        tf.write(data.tobytes())
        tf.flush()
        argv = ["test_tx_vectors.py", "-f", tf.name, "-b", "2"]
        import socket as _socket
        with mock.patch.object(sys, "argv", argv), mock.patch.object(_socket, "socket", Recorder):
            runpy.run_path(os.path.join(REF, "pipeline/test_transmitters/test_tx_vectors.py"), run_name="__main__")
    assert len(sent) == npkt and len(set(len(p) for p in sent)) == 1
    np.savez_compressed(os.path.join(HERE, "golden_snap2_packets.npz"), data=data,
                        packets=np.frombuffer(b"".join(sent), dtype=np.uint8).reshape(npkt, -1), nchan_blocks=2,
                        nstand_per_pkt=32)
    print("snap2", npkt, "packets of", len(sent[0]), "bytes")


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "snap2":
        snap2_packets()
    else:
        main()
        snap2_packets()
