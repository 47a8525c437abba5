"""ctypes binding of libvcmts_probe.so (include/vcmts_b200_probe.h): hardware probes used by
scripts/ and one GPU test.  Diagnostics only -- the product (libvcmts_b200.so) does not load it."""
import ctypes
import os
from ctypes import c_int, c_int32, c_void_p

HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(HERE, "libvcmts_probe.so")
_P = c_void_p
_PROTOS = {
    "vcm_umma_probe": (c_int, [c_int32, c_int32, c_int32, _P, _P]),
    "vcm_umma_rate": (c_int, [c_int32, c_int32, _P, _P]),
    "vcm_umma_rate2": (c_int, [c_int32, c_int32, c_int32, c_int32, _P, _P]),
    "vcm_umma_rate3": (c_int, [c_int32, c_int32, c_int32, _P, _P]),
    "vcm_tma_probe": (c_int, [_P, c_int32, c_int32, c_int32, c_int32, c_int32, c_int32, _P, _P]),
    "vcm_umma_acc_probe": (c_int, [_P, _P, _P, _P]),
}
_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(SO):
            from . import build as _build
            _build.build_probe()
        L = ctypes.CDLL(SO)
        for name, (res, args) in _PROTOS.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib
