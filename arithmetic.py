"""Drop-in for /root/reference/arithmetic.py -- same five signatures, B200 kernels underneath.

    vec_add(a, b, MOD)   arithmetic.py:3-5     (a + b) % MOD
    vec_sub(a, b, MOD)   arithmetic.py:7-9     (a - b) % MOD   (non-negative, Python semantics)
    vec_mul(a, b, MOD)   arithmetic.py:11-13   (a * b) % MOD   (exact: the reference overflows on
                                               int64 arrays for >31-bit moduli; this does not)
    NTT(x)               arithmetic.py:15-16   identity stub in the reference; here the forward
                                               negacyclic NTT, NTT(x)[k] = x(psi^(2*brv(k)+1))
    iNTT(x)              arithmetic.py:18-19   identity stub in the reference; here its inverse

Operands: torch CUDA tensors (int64/uint64, layout [L, B, N], N contiguous) are used in place;
numpy arrays / CPU tensors are copied to the GPU and the result comes back in the same container.
Out-of-place like the reference: inputs are never modified.  For vec_* `MOD` is anything numpy
broadcasts against the operands (a scalar, one modulus per limb as [L,1,1], per batch item, per
coefficient ...); the result has the broadcast shape.  NTT/iNTT take their moduli from the
keyword-only `MOD=` or from the ring installed with primitive.set_ring / `with primitive.ring(...)`.

Input ranges differ between the two families:
  * vec_add / vec_sub / vec_mul accept ANY integer that fits the container (negative int64 values,
    uint64 values up to 2^64-1, object-dtype Python ints that fit 64 bits) and return canonical
    residues in [0, MOD) -- Python's `%` semantics, like the reference on dtype=object.
  * NTT / iNTT (and poly_mul, poly_mul_pre, poly_mul_ntt, poly_mac_ntt, automorphism in polynomial.py)
    require canonical residues in [0, q): the kernels' lazy-reduction schedule starts from q.
    Centred or negative coefficients must be reduced first, e.g. x = vec_add(x, 0 * x, MOD).
    GFHE_CHECK_RANGE=1 in the environment turns on a (slow) check that raises ValueError instead of
    returning wrong residues.
No CPU fallback.
"""
import importlib.util as _ilu
import os as _os
import sys as _sys

import numpy as np  # noqa: F401  (the reference leaks `np` through its star-import chain)


def _load_pkg():
    name = "gpu_fhe_b200"
    if name in _sys.modules:
        return _sys.modules[name]
    root = _os.path.join(_os.path.dirname(_os.path.abspath(__file__)), "gpu-fhe_b200")
    spec = _ilu.spec_from_file_location(name, _os.path.join(root, "__init__.py"), submodule_search_locations=[root])
    mod = _ilu.module_from_spec(spec)
    _sys.modules[name] = mod
    try:
        spec.loader.exec_module(mod)
    except BaseException:
        del _sys.modules[name]
        raise
    return mod


_pkg = _load_pkg()
_d = _pkg.dispatch

__all__ = ["np", "vec_add", "vec_sub", "vec_mul", "NTT", "iNTT"]


def vec_add(a, b, MOD):
    assert a.shape == b.shape
    return _d.vec_op(0, a, b, MOD)


def vec_sub(a, b, MOD):
    assert a.shape == b.shape
    return _d.vec_op(1, a, b, MOD)


def vec_mul(a, b, MOD):
    assert a.shape == b.shape
    return _d.vec_op(2, a, b, MOD)


def NTT(x, *, MOD=None):
    return _d.ntt(x, MOD, inverse=False)


def iNTT(x, *, MOD=None):
    return _d.ntt(x, MOD, inverse=True)
