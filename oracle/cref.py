"""ctypes binding of oracle/ntt_oracle.c.  TEST INFRASTRUCTURE ONLY (see oracle.py header).

Used by tests/ as the full-size checker and by bench.py as the timed CPU baseline.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "libntt_oracle.so")
_lib = None


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "ntt_oracle.c")
    if force or not os.path.exists(_LIB) or os.path.getmtime(_LIB) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s", "-B", "libntt_oracle.so"])
    return _LIB


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB):
            build()
        L = ctypes.CDLL(_LIB)
        u64p = ctypes.POINTER(ctypes.c_uint64)
        L.oracle_ctx_create.restype = ctypes.c_void_p
        L.oracle_ctx_create.argtypes = [ctypes.c_uint32, ctypes.c_uint64, ctypes.c_uint64]
        L.oracle_ctx_destroy.argtypes = [ctypes.c_void_p]
        for name in ("oracle_ntt_fwd_exact", "oracle_ntt_inv_exact", "oracle_ntt_fwd_fast", "oracle_ntt_inv_fast"):
            getattr(L, name).argtypes = [ctypes.c_void_p, u64p]
            getattr(L, name).restype = None
        L.oracle_poly_mul.argtypes = [ctypes.c_void_p, u64p, u64p, u64p, ctypes.c_int]
        L.oracle_poly_mul.restype = None
        L.oracle_schoolbook.argtypes = [u64p, u64p, u64p, ctypes.c_size_t, ctypes.c_uint64]
        L.oracle_schoolbook.restype = None
        for name in ("oracle_vec_add", "oracle_vec_sub", "oracle_vec_mul"):
            getattr(L, name).argtypes = [u64p, u64p, u64p, ctypes.c_size_t, ctypes.c_uint64]
            getattr(L, name).restype = None
        L.oracle_batch.argtypes = [ctypes.POINTER(ctypes.c_void_p), ctypes.c_int, ctypes.c_int, u64p, u64p, u64p,
                                   ctypes.c_uint32, ctypes.c_uint32, ctypes.c_uint32, ctypes.c_int]
        L.oracle_batch.restype = ctypes.c_int
        _lib = L
    return _lib


def _p(a: np.ndarray):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_uint64))


def _u64(a) -> np.ndarray:
    return np.ascontiguousarray(np.asarray(a).astype(np.uint64, copy=False))


class Ring:
    """Per-limb oracle contexts for an RNS ring (moduli, psis) at degree N."""

    def __init__(self, N: int, moduli, psis):
        self.N = int(N)
        self.logn = self.N.bit_length() - 1
        assert 1 << self.logn == self.N
        self.moduli = [int(q) for q in moduli]
        self.psis = [int(p) for p in psis]
        self._ctx = [lib().oracle_ctx_create(self.logn, q, p) for q, p in zip(self.moduli, self.psis)]
        self._arr = (ctypes.c_void_p * len(self._ctx))(*self._ctx)

    def __del__(self):
        try:
            for c in self._ctx:
                lib().oracle_ctx_destroy(c)
        except Exception:
            pass

    def _batch(self, op, a, b, fast, nthreads):
        a = _u64(a)
        assert a.ndim == 3 and a.shape[0] == len(self._ctx) and a.shape[2] == self.N, a.shape
        out = np.empty_like(a)
        bb = _u64(b) if b is not None else a
        lib().oracle_batch(self._arr, op, int(fast), _p(a), _p(bb), _p(out), self.logn, a.shape[0], a.shape[1], int(nthreads))
        return out

    def ntt(self, x, fast=False, nthreads=1):
        return self._batch(0, x, None, fast, nthreads)

    def intt(self, x, fast=False, nthreads=1):
        return self._batch(1, x, None, fast, nthreads)

    def poly_mul(self, a, b, fast=False, nthreads=1):
        return self._batch(2, a, b, fast, nthreads)


def schoolbook(a, b, q: int) -> np.ndarray:
    a, b = _u64(a), _u64(b)
    c = np.empty_like(a)
    lib().oracle_schoolbook(_p(a), _p(b), _p(c), a.size, int(q))
    return c


def vec_op(name: str, a, b, q: int) -> np.ndarray:
    a, b = _u64(a), _u64(b)
    assert a.shape == b.shape
    out = np.empty_like(a)
    getattr(lib(), "oracle_vec_" + name)(_p(a), _p(b), _p(out), a.size, int(q))
    return out
