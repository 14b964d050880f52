"""Regenerates the golden fixtures in this directory.  Run in the BUILD container only:

    python tests/golden/make_golden.py

* vec_ops.npz   -- outputs of the REFERENCE's own vec_add / vec_sub / vec_mul
                   (/root/reference/arithmetic.py:3-13), imported from /root/reference and run on
                   dtype=object arrays (its only exact mode) -- this pins the element-wise ops.
                   Also records that the reference's NTT/iNTT are identity stubs.
* ntt_small.npz -- NTT(x)[k] = x(psi^(2*brv(k)+1)) evaluated straight from the definition
                   (oracle.ntt_definition, O(N^2) Horner) -- the reference has no NTT to pin to.
* polymul.npz   -- negacyclic products by schoolbook convolution and by Kronecker substitution
                   (neither uses roots of unity or an NTT).

/root/reference does not exist on the GPU box; tests only read the .npz files.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import oracle as O  # noqa: E402


def load_reference():
    import importlib.util
    sys.path.insert(0, "/root/reference")
    import arithmetic as ref_arith  # the reference module itself
    assert ref_arith.__file__.startswith("/root/reference"), ref_arith.__file__
    spec = importlib.util.spec_from_file_location("ref_polynomial", "/root/reference/ polynomial.py")
    ref_poly = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref_poly)
    sys.path.pop(0)
    return ref_arith, ref_poly


def gen_vec(ref):
    out = {}
    rng = np.random.default_rng(20261104)
    cases = []
    # (name, moduli list (per leading index) or scalar, shape, value range)
    q30, q50, q60 = O.ntt_primes(30, 1)[0], O.ntt_primes(50, 4), O.ntt_primes(60, 3)
    cases.append(("c1_30bit", [q30], (1, 1, 1024), "canon"))
    cases.append(("c2_50bit", q50, (4, 3, 257), "canon"))          # odd inner length: scalar tail path
    cases.append(("c4_60bit", q60, (3, 2, 512), "canon"))
    cases.append(("anyint64", [q60[0]], (1, 1, 600), "any"))       # arbitrary int64 incl. negatives
    cases.append(("smallmod", [7], (1, 1, 64), "any"))
    cases.append(("evenmod", [1 << 40], (1, 1, 64), "any"))        # reference accepts any MOD >= 1
    names = []
    for name, moduli, shape, mode in cases:
        L = shape[0]
        a = np.empty(shape, dtype=np.int64)
        b = np.empty(shape, dtype=np.int64)
        for l in range(L):
            if mode == "canon":
                a[l] = rng.integers(0, moduli[l], shape[1:], dtype=np.int64)
                b[l] = rng.integers(0, moduli[l], shape[1:], dtype=np.int64)
            else:
                a[l] = rng.integers(-2**63, 2**63 - 1, shape[1:], dtype=np.int64)
                b[l] = rng.integers(-2**63, 2**63 - 1, shape[1:], dtype=np.int64)
                a[l].flat[:4] = [0, -1, 2**63 - 1, -2**63]
                b[l].flat[:4] = [-2**63, 2**63 - 1, -1, 0]
        MOD = np.array(moduli, dtype=object).reshape((L,) + (1,) * (len(shape) - 1))
        ao, bo = a.astype(object), b.astype(object)
        out[name + "_a"], out[name + "_b"] = a, b
        out[name + "_mod"] = np.array(moduli, dtype=np.uint64)
        for op in ("vec_add", "vec_sub", "vec_mul"):
            r = getattr(ref, op)(ao, bo, MOD)            # <- the reference, dtype=object
            assert all(0 <= int(v) < moduli[i // (r.size // L)] for i, v in enumerate(r.ravel()))
            out[name + "_" + op] = r.astype(np.uint64)
        names.append(name)
    out["names"] = np.array(names)
    # the reference's transforms are identity stubs: record it so a test documents "parity unpinned"
    x = np.arange(8)
    out["ref_ntt_is_identity"] = np.array([ref.NTT(x) is x and ref.iNTT(x) is x])
    return out


def gen_ntt():
    out, names = {}, []
    import random
    rng = random.Random(20261105)
    for N, bits in [(16, 30), (16, 60), (64, 50), (256, 60), (1024, 30), (1024, 60)]:
        q = O.ntt_primes(bits, 1)[0]
        psi = O.find_psi(q, N)
        x = O.random_poly(rng, N, q)
        x[0], x[1] = 0, q - 1
        X = O.ntt_definition(x, q, psi)
        assert O.intt_definition(X, q, psi) == x
        name = "n%d_b%d" % (N, bits)
        out[name + "_q"] = np.array([q, psi], dtype=np.uint64)
        out[name + "_x"] = np.array(x, dtype=np.uint64)
        out[name + "_X"] = np.array(X, dtype=np.uint64)
        names.append(name)
    out["names"] = np.array(names)
    return out


def gen_polymul():
    out, names = {}, []
    import random
    rng = random.Random(20261106)
    for N, bits, how in [(16, 60, "schoolbook"), (256, 60, "schoolbook"), (1024, 30, "schoolbook"),
                         (4096, 50, "kronecker"), (16384, 60, "kronecker")]:
        q = O.ntt_primes(bits, 1)[0]
        a, b = O.random_poly(rng, N, q), O.random_poly(rng, N, q)
        c = O.schoolbook_negacyclic(a, b, q) if how == "schoolbook" else O.kronecker_negacyclic(a, b, q)
        if N <= 4096:
            assert c == O.kronecker_negacyclic(a, b, q)
        name = "n%d_b%d" % (N, bits)
        out[name + "_q"] = np.array([q], dtype=np.uint64)
        out[name + "_a"] = np.array(a, dtype=np.uint64)
        out[name + "_b"] = np.array(b, dtype=np.uint64)
        out[name + "_c"] = np.array(c, dtype=np.uint64)
        names.append(name)
    out["names"] = np.array(names)
    return out


if __name__ == "__main__":
    ref_arith, ref_poly = load_reference()
    np.savez_compressed(os.path.join(HERE, "vec_ops.npz"), **gen_vec(ref_arith))
    np.savez_compressed(os.path.join(HERE, "ntt_small.npz"), **gen_ntt())
    np.savez_compressed(os.path.join(HERE, "polymul.npz"), **gen_polymul())
    # " polynomial.py":3-5 returns None: recorded in the test-suite as a documented deviation
    assert ref_poly.poly_add((np.arange(3), np.arange(3)), (np.arange(3), np.arange(3)), 5) is None
    for f in ("vec_ops.npz", "ntt_small.npz", "polymul.npz"):
        print(f, os.path.getsize(os.path.join(HERE, f)), "bytes")
