import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "caltech-bifrost-dsp_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN


@pytest.fixture
def host_only(monkeypatch):
    """Run the blocks' host logic without a GPU: libxeng entry points are replaced by the tag-moving doubles of
    tests/fakes.py and 'cuda' allocations land in host memory.  Only for `-m "not gpu"` tests of state machines,
    headers and command handling; the kernels themselves are covered by the gpu tests."""
    import fakes
    from xengine import ops, ringshim
    from xengine import runtime as rt

    real_alloc = ringshim.alloc_bytes

    def host_alloc(nbytes, space):
        return real_alloc(nbytes, "system" if space == "cuda" else space)

    monkeypatch.setattr(ringshim, "alloc_bytes", host_alloc)
    monkeypatch.setattr(rt, "alloc_bytes", host_alloc)
    monkeypatch.setattr(rt, "set_device", lambda gpu: None)
    monkeypatch.setattr(ops, "XCorr", fakes.FakeXCorr)
    monkeypatch.setattr(ops, "acc", fakes.fake_acc)
    monkeypatch.setattr(ops, "subsel", fakes.fake_subsel)
    monkeypatch.setattr(ops, "reorder_packets", fakes.fake_reorder_packets)
    monkeypatch.setattr(ops, "Beamformer", fakes.FakeBeamformer)
    del fakes.CALLS[:]
    return fakes
