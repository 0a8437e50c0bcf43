"""ctypes binding of libpcg_host.so (include/pcg_host.h): host range coder + CDF-table builder (csrc/host_rans.cpp)
and the latent-coordinate octree coder (csrc/host_octree.cpp).

The coder stays on the host by design (north-star); it consumes the int32 symbol / index buffers the
GPU kernels produce (one pinned D2H copy instead of the reference's `.tolist()`).
"""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libpcg_host.so")
_lib = None


def load():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} is missing: run __graft_entry__.build()")
        lib = ctypes.CDLL(LIB_PATH)
        lib.pcg_rans_encode.restype = ctypes.c_long
        lib.pcg_rans_encode.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_long, ctypes.c_void_p, ctypes.c_int,
                                        ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_long]
        lib.pcg_rans_decode.restype = ctypes.c_int
        lib.pcg_rans_decode.argtypes = [ctypes.c_void_p, ctypes.c_long, ctypes.c_void_p, ctypes.c_long, ctypes.c_void_p,
                                        ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
        lib.pcg_pmf_to_quantized_cdf.restype = ctypes.c_int
        lib.pcg_pmf_to_quantized_cdf.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
        for name in ("pcg_octree_encode", "pcg_octree_decode", "pcg_octree_count"):
            getattr(lib, name).restype = ctypes.c_long
        lib.pcg_octree_encode.argtypes = [ctypes.c_void_p, ctypes.c_long, ctypes.c_int, ctypes.c_void_p, ctypes.c_long]
        lib.pcg_octree_count.argtypes = [ctypes.c_void_p, ctypes.c_long]
        lib.pcg_octree_decode.argtypes = [ctypes.c_void_p, ctypes.c_long, ctypes.c_void_p, ctypes.c_long]
        _lib = lib
    return _lib


_LONG_MIN = -(1 << 63)


def octree_encode(xyz, stride=1):
    """Lossless bytes of a SET of lattice points (csrc/host_octree.cpp): xyz (n,3) integers on a `stride` lattice.
    Stands where the reference shells out to tmc3 (model/model.py:388-431 `gpcc_encode`)."""
    xyz = _i32(np.asarray(xyz).reshape(-1, 3))
    cap = xyz.shape[0] * 2 + 256
    while True:
        out = np.empty(cap, dtype=np.uint8)
        nb = load().pcg_octree_encode(xyz.ctypes.data, xyz.shape[0], int(stride), out.ctypes.data, cap)
        if nb == _LONG_MIN:
            raise ValueError(f"octree_encode: points off the stride-{stride} lattice or extent beyond 21 bits per axis")
        if nb >= 0:
            return out[:nb].tobytes()
        cap = -nb


def octree_decode(data):
    """(n,3) int32 points of an `octree_encode` stream, in Morton order (model/model.py:445-486 `gpcc_decode`)."""
    buf = np.frombuffer(data, dtype=np.uint8)
    n = load().pcg_octree_count(buf.ctypes.data, buf.size)
    if n == _LONG_MIN:
        raise ValueError("octree_decode: not a latent-coordinate stream")
    out = np.empty((max(n, 1), 3), dtype=np.int32)
    got = load().pcg_octree_decode(buf.ctypes.data, buf.size, out.ctypes.data, n)
    if got != n:
        raise ValueError("octree_decode: corrupt latent-coordinate stream")
    return out[:n]


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def encode(symbols, indexes, cdfs, cdf_sizes, offsets):
    symbols, indexes, cdfs, cdf_sizes, offsets = map(_i32, (symbols, indexes, cdfs, cdf_sizes, offsets))
    n = symbols.size
    cap = n * 4 + 1024
    while True:
        out = np.empty(cap, dtype=np.uint8)
        nb = load().pcg_rans_encode(symbols.ctypes.data, indexes.ctypes.data, n, cdfs.ctypes.data, cdfs.shape[1],
                                    cdf_sizes.ctypes.data, offsets.ctypes.data, out.ctypes.data, cap)
        if nb >= 0:
            return out[:nb].tobytes()
        cap = -nb


def decode(encoded, indexes, cdfs, cdf_sizes, offsets):
    indexes, cdfs, cdf_sizes, offsets = map(_i32, (indexes, cdfs, cdf_sizes, offsets))
    enc = np.frombuffer(encoded, dtype=np.uint8)
    out = np.empty(indexes.size, dtype=np.int32)
    rc = load().pcg_rans_decode(enc.ctypes.data, enc.size, indexes.ctypes.data, indexes.size, cdfs.ctypes.data,
                                cdfs.shape[1], cdf_sizes.ctypes.data, offsets.ctypes.data, out.ctypes.data)
    if rc != 0:
        raise ValueError("corrupt range-coder stream")
    return out


def pmf_to_quantized_cdf(pmf, precision=16):
    pmf = np.ascontiguousarray(pmf, dtype=np.float32)
    out = np.empty(pmf.size + 1, dtype=np.int32)
    rc = load().pcg_pmf_to_quantized_cdf(pmf.ctypes.data, pmf.size, precision, out.ctypes.data)
    if rc == 1:
        raise ValueError("Invalid `pmf`, non-finite or negative element found")
    if rc == 2:
        raise ValueError("Invalid `pmf`: at least one element must have a non-zero probability.")
    if rc != 0:
        raise ValueError("pmf_to_quantized_cdf failed")
    return out
