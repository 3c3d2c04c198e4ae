"""ctypes binding of libxeng.so (include/xeng.h) -- the stand-in for ``from bifrost.libbifrost import _bf``
on the X-engine hot path (reference call sites: corr_block.py:253,331,445; corr_subsel_block.py:298;
corr_output_full_block.py:669; beamform_block.py:251,449; beamform_sum_beams_block.py:245).

There is no CPU fallback: if the shared library is missing or a call fails, an exception is raised.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("XENG_LIB_PATH") or os.path.join(_HERE, "libxeng.so")   # override: developer timing builds

XENG_OK = 0


class XengError(RuntimeError):
    pass


_lib = None

_vp, _i, _i64, _sz, _u64 = C.c_void_p, C.c_int, C.c_int64, C.c_size_t, C.c_uint64

# name -> (restype, argtypes); must list every symbol include/xeng.h declares
SIGNATURES = {
    "xeng_last_error": (C.c_char_p, []),
    "xeng_version": (_i, []),
    "xeng_device_count": (_i, []),
    "xeng_corr_create": (_i, [C.POINTER(_vp), _i, _i, _i, _i, _i]),
    "xeng_corr_destroy": (_i, [_vp]),
    "xeng_corr_set_stream": (_i, [_vp, _vp]),
    "xeng_corr_set_impl": (_i, [_vp, _i]),
    "xeng_corr_matlen": (_i64, [_vp]),
    "xeng_corr_run": (_i, [_vp, _vp, _vp, _i]),
    "xeng_corr_run_host": (_i, [_vp, _vp, _vp, _i]),
    "xeng_corr_accumulate": (_i, [_vp, _vp, _vp, _i]),
    "xeng_corr_set_subsel": (_i, [_vp, _vp, _vp, _i, _i]),
    "xeng_corr_accumulate_subsel": (_i, [_vp, _vp, _vp, _i, _vp, _i]),
    "xeng_corr_sync": (_i, [_vp]),
    "xeng_get_order": (_i, [_vp, _vp, _vp, _i, _i]),
    "xeng_acc": (_i, [_vp, _vp, _sz, _i, _vp]),
    "xeng_subsel": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i64, _vp]),
    "xeng_reorder": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i64, _vp]),
    "xeng_snap2_assemble": (_i, [_vp, _i, _i64, _vp, _u64, _i, _i, _i, _i, _i, _vp, _vp]),
    "xeng_upchan_fft": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _vp]),
    "xeng_upchan_beamform": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _i, _vp]),
    "xeng_upchan_correlate": (_i, [_vp, _vp, _i, _i, _i, _i, _vp]),
    "xeng_subsel_packets": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i64, _i, _vp]),
    "xeng_reorder_packets": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i64, _i, _vp]),
    "xeng_memcpy_async": (_i, [_vp, _vp, _sz, _vp]),
    "xeng_beam_power_static": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _i, _vp]),
    "xeng_peer_push": (_i, [_vp, _vp, _sz, _vp, C.c_longlong, _vp]),
    "xeng_peer_move": (_i, [_vp, _vp, _sz, _vp, C.c_longlong, _vp]),
    "xeng_bf_create": (_i, [C.POINTER(_vp), _i, _i, _i, _i, _i, _i]),
    "xeng_bf_destroy": (_i, [_vp]),
    "xeng_bf_set_stream": (_i, [_vp, _vp]),
    "xeng_bf_run": (_i, [_vp, _vp, _vp, _vp]),
    "xeng_bf_set_weights": (_i, [_vp, _vp]),
    "xeng_bf_set_impl": (_i, [_vp, _i]),
    "xeng_bf_integrate": (_i, [_vp, _vp, _vp, _i]),
    "xeng_bf_integrate_single": (_i, [_vp, _vp, _vp, _i, _i]),
    "xeng_bf_run_integrate": (_i, [_vp, _vp, _vp, _vp, _i]),
    "xeng_bf_voltages": (_vp, [_vp]),
    "xeng_bf_set_tap": (_i, [_vp, _i]),
    "xeng_bf_vlbi_tap": (_i, [_vp, _vp, _vp, _i]),
    "xeng_bf_sync": (_i, [_vp]),
}


def load():
    """Load libxeng.so (once).  Raises XengError when it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise XengError(
                "libxeng.so not found at %s: build it with `make -C %s` (or __graft_entry__.build()); "
                "there is no CPU fallback" % (LIB_PATH, os.path.join(_HERE, "csrc")))
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def last_error():
    return load().xeng_last_error().decode("utf-8", "replace")


def check(status, what=""):
    """Mirror of the reference's `if status != _bf.BF_STATUS_SUCCESS: raise RuntimeError`
    (corr_block.py:254-256, corr_subsel_block.py:299-301)."""
    if status != XENG_OK:
        raise XengError("%s failed with status %d: %s" % (what or "libxeng call", status, last_error()))
    return status


def ptr(x):
    """Raw address of a buffer: torch tensor (device or host), numpy array, int, bifrost ndarray or
    anything exposing ``.data_ptr()`` / ``.ctypes.data`` / ``__cuda_array_interface__``."""
    if x is None:
        return None
    if isinstance(x, int):
        return x
    if hasattr(x, "cuda_stream"):      # torch.cuda.Stream
        return x.cuda_stream
    if hasattr(x, "data_ptr"):
        return x.data_ptr()
    cai = getattr(x, "__cuda_array_interface__", None)
    if cai is not None:
        return cai["data"][0]
    if hasattr(x, "ctypes"):
        return x.ctypes.data
    raise TypeError("cannot take the address of %r" % type(x))
