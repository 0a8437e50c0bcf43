"""ctypes front for oracle/csrc/sparse_conv_ref.cpp: the C++ restatement of ME's CPU path
(hash map -> kernel map -> per-offset gather / OpenBLAS sgemm / scatter-add).
TEST INFRASTRUCTURE and bench.py's timed CPU baseline."""
import ctypes
import glob
import os

import numpy as np

from .build import lib_path

_lib = None
_blas = None


def _find_openblas():
    import scipy
    cands = glob.glob(os.path.join(os.path.dirname(scipy.__file__), "..", "scipy.libs", "libscipy_openblas*.so"))
    if not cands:
        raise RuntimeError("scipy's bundled OpenBLAS not found")
    return cands[0]


def L():
    global _lib, _blas
    if _lib is None:
        _blas = ctypes.CDLL(_find_openblas())
        _lib = ctypes.CDLL(lib_path("libsparse_conv_ref.so"))
        _lib.ref_set_sgemm(ctypes.cast(_blas.scipy_cblas_sgemm, ctypes.c_void_p))
        _lib.ref_stride_coords.restype = ctypes.c_long
        _lib.ref_generate_coords.restype = ctypes.c_long
        set_threads(os.cpu_count() or 1)
    return _lib


def set_threads(n):
    if _lib is None:
        L()
    # OpenMP owns the cores (row blocks); the per-offset GEMMs are too small for BLAS-level threading and the
    # two thread pools otherwise fight over the cores (measured: 100x slowdown)
    _blas.scipy_openblas_set_num_threads(1)
    _lib.ref_set_threads(int(n))
    return int(n)


def _p(a):
    return ctypes.c_void_p(a.ctypes.data)


def stride_coords(C, s):
    C = np.ascontiguousarray(C, np.int32)
    out = np.empty_like(C)
    n = L().ref_stride_coords(_p(C), ctypes.c_long(len(C)), int(s), _p(out))
    return out[:n].copy()


def generate_coords(C, ksize, ts_out):
    C = np.ascontiguousarray(C, np.int32)
    out = np.empty((len(C) * ksize ** 3, 4), np.int32)
    n = L().ref_generate_coords(_p(C), ctypes.c_long(len(C)), int(ksize), int(ts_out), _p(out))
    return out[:n].copy()


def kernel_map(Cin, Cout, ksize, ts, transposed=False):
    Cin = np.ascontiguousarray(Cin, np.int32)
    Cout = np.ascontiguousarray(Cout, np.int32)
    nbr = np.empty((ksize ** 3, len(Cout)), np.int32)
    L().ref_kernel_map(_p(Cin), ctypes.c_long(len(Cin)), _p(Cout), ctypes.c_long(len(Cout)), int(ksize), int(ts),
                       int(bool(transposed)), _p(nbr))
    return nbr


def conv(F, nbr, W, bias=None, act=0):
    F = np.ascontiguousarray(F, np.float32)
    W = np.ascontiguousarray(W, np.float32)
    if W.ndim == 2:
        W = W[None]
    K, cin, cout = W.shape
    assert nbr.shape[0] == K and F.shape[1] == cin
    n_out = nbr.shape[1]
    out = np.empty((n_out, cout), np.float32)
    b = np.ascontiguousarray(bias, np.float32).reshape(-1) if bias is not None else None
    rc = L().ref_conv(_p(F), ctypes.c_long(len(F)), cin, _p(nbr), K, ctypes.c_long(n_out), _p(W),
                      _p(b) if b is not None else None, cout, int(act), _p(out))
    assert rc == 0
    return out


def gdn(X, beta, gamma, inverse):
    X = np.ascontiguousarray(X, np.float32)
    beta = np.ascontiguousarray(beta, np.float32)
    gamma = np.ascontiguousarray(gamma, np.float32)
    out = np.empty_like(X)
    rc = L().ref_gdn(_p(X), ctypes.c_long(len(X)), X.shape[1], _p(beta), _p(gamma), int(bool(inverse)), _p(out))
    assert rc == 0
    return out
