"""ctypes front for oracle/csrc/rans64_ref.c (compressai's rANS coder restated).  TEST INFRASTRUCTURE."""
import ctypes

import numpy as np

from .build import lib_path

_lib = None


def _L():
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(lib_path("librans64_ref.so"))
        _lib.rans_encode_with_indexes.restype = ctypes.c_long
        _lib.rans_encode_with_indexes.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_long, ctypes.c_void_p,
                                                  ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_long]
        _lib.rans_decode_with_indexes.restype = None
        _lib.rans_decode_with_indexes.argtypes = [ctypes.c_char_p, ctypes.c_void_p, ctypes.c_long, ctypes.c_void_p,
                                                  ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
    return _lib


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def encode_with_indexes(symbols, indexes, cdfs, cdf_sizes, offsets):
    symbols, indexes, cdfs, cdf_sizes, offsets = map(_i32, (symbols, indexes, cdfs, cdf_sizes, offsets))
    n = symbols.size
    cap = n * 48 + 64
    out = np.empty(cap, dtype=np.uint8)
    nb = _L().rans_encode_with_indexes(symbols.ctypes.data, indexes.ctypes.data, n, cdfs.ctypes.data, cdfs.shape[1],
                                       cdf_sizes.ctypes.data, offsets.ctypes.data, out.ctypes.data, cap)
    assert nb >= 0
    return out[:nb].tobytes()


def decode_with_indexes(encoded, indexes, cdfs, cdf_sizes, offsets):
    indexes, cdfs, cdf_sizes, offsets = map(_i32, (indexes, cdfs, cdf_sizes, offsets))
    out = np.empty(indexes.size, dtype=np.int32)
    _L().rans_decode_with_indexes(encoded, indexes.ctypes.data, indexes.size, cdfs.ctypes.data, cdfs.shape[1],
                                  cdf_sizes.ctypes.data, offsets.ctypes.data, out.ctypes.data)
    return out
