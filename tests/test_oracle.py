"""CPU suite, part 1: pin the oracle.

* vec_* restatements against the REFERENCE's outputs (tests/golden/vec_ops.npz, produced by
  importing /root/reference/arithmetic.py -- see tests/golden/make_golden.py).
* NTT / iNTT / poly-mul: the reference has none (identity stubs), so the oracle is pinned to the
  mathematical definition: definition-level evaluation, schoolbook and Kronecker products, and the
  C oracle's two arithmetic flavours must all agree.
"""
import random

import numpy as np
import pytest

from oracle import cref
from oracle import oracle as O


@pytest.fixture(scope="module")
def vec_golden(golden_dir):
    return np.load(golden_dir + "/vec_ops.npz")


def test_reference_transforms_are_identity_stubs(vec_golden):
    # documents why NTT parity is "unpinned": arithmetic.py:15-19 return their argument
    assert bool(vec_golden["ref_ntt_is_identity"][0])


@pytest.mark.parametrize("op", ["vec_add", "vec_sub", "vec_mul"])
def test_vec_oracles_match_reference_outputs(vec_golden, op):
    for name in vec_golden["names"]:
        a, b, mod = vec_golden[name + "_a"], vec_golden[name + "_b"], vec_golden[name + "_mod"]
        want = vec_golden[name + "_" + op]
        for l in range(a.shape[0]):
            q = int(mod[l])
            got_py = getattr(O, op)(a[l].ravel().tolist(), b[l].ravel().tolist(), q)
            assert got_py == want[l].ravel().tolist(), (name, l)
            if a[l].min() >= 0:   # the C oracle takes unsigned words
                got_c = cref.vec_op(op[4:], a[l].ravel(), b[l].ravel(), q)
                assert np.array_equal(got_c, want[l].ravel()), (name, l)


def test_root_convention_pinned_values():
    # SURVEY.md section 8c
    assert O.find_psi(1073479681, 1024) == 530309095
    assert O.find_psi(1125899902124033, 4096) == 93790016731446
    assert O.find_psi(1152921504606584833, 65536) == 987813353222176621
    assert O.find_psi(1152921504606584833, 131072) == 720645352895426071
    p60 = O.ntt_primes(60, 64)
    assert p60[0] == 1152921504606584833 and p60[63] == 1152921504241942529 and len(set(p60)) == 64
    assert O.ntt_primes(50, 4) == [1125899902124033, 1125899886395393, 1125899884036097, 1125899882987521]
    assert O.ntt_primes(30, 2) == [1073479681, 1068236801]


def test_ntt_golden_matches_all_oracles(golden_dir):
    g = np.load(golden_dir + "/ntt_small.npz")
    for name in g["names"]:
        q, psi = (int(v) for v in g[name + "_q"])
        x, X = g[name + "_x"], g[name + "_X"]
        N = len(x)
        assert O.ntt_fast(x.tolist(), q, psi) == X.tolist()
        assert O.intt_fast(X.tolist(), q, psi) == x.tolist()
        R = cref.Ring(N, [q], [psi])
        for fast in (False, True):
            assert np.array_equal(R.ntt(x.reshape(1, 1, N), fast=fast).ravel(), X)
            assert np.array_equal(R.intt(X.reshape(1, 1, N), fast=fast).ravel(), x)


def test_polymul_golden_matches_all_oracles(golden_dir):
    g = np.load(golden_dir + "/polymul.npz")
    for name in g["names"]:
        q = int(g[name + "_q"][0])
        a, b, c = g[name + "_a"], g[name + "_b"], g[name + "_c"]
        N = len(a)
        psi = O.find_psi(q, N)
        R = cref.Ring(N, [q], [psi])
        for fast in (False, True):
            assert np.array_equal(R.poly_mul(a.reshape(1, 1, N), b.reshape(1, 1, N), fast=fast).ravel(), c), name
        if N <= 4096:
            assert O.poly_mul(a.tolist(), b.tolist(), q, psi) == c.tolist()
        if N <= 1024:
            assert np.array_equal(cref.schoolbook(a, b, q), c)
        # the pair of cheap spot checks used at full size must accept the true product ...
        rng = random.Random(3)
        for _ in range(3):
            k = rng.randrange(N)
            assert O.check_point_eval(a.tolist(), b.tolist(), c.tolist(), q, psi, k)
            assert O.schoolbook_coeff(a.tolist(), b.tolist(), q, k) == int(c[k])
        # ... and reject a corrupted one
        bad = c.copy()
        bad[5] = (int(bad[5]) + 1) % q
        assert O.schoolbook_coeff(a.tolist(), b.tolist(), q, 5) != int(bad[5])
        assert not O.check_point_eval(a.tolist(), b.tolist(), bad.tolist(), q, psi, 1)


@pytest.mark.parametrize("logn,bits", [(4, 30), (7, 50), (10, 60), (13, 60)])
def test_c_oracle_properties(logn, bits):
    N = 1 << logn
    qs = O.ntt_primes(bits, 2)
    psis = [O.find_psi(q, N) for q in qs]
    R = cref.Ring(N, qs, psis)
    rng = np.random.default_rng(logn)
    a = np.stack([rng.integers(0, q, (3, N), dtype=np.uint64) for q in qs])
    b = np.stack([rng.integers(0, q, (3, N), dtype=np.uint64) for q in qs])
    A = R.ntt(a)
    assert np.array_equal(A, R.ntt(a, fast=True, nthreads=3))
    assert np.array_equal(R.intt(A), a) and np.array_equal(R.intt(A, fast=True), a)
    # evaluation order: NTT(x)[k] = x(psi^(2 brv(k)+1))
    for l in range(2):
        for k in (0, 1, N - 1):
            assert int(A[l, 0, k]) == O.ntt_point(a[l, 0].tolist(), qs[l], psis[l], k)
    c = R.poly_mul(a, b)
    assert np.array_equal(c, R.poly_mul(a, b, fast=True, nthreads=2))
    # multiplication by X: negacyclic shift with sign flip
    xpoly = np.zeros_like(a)
    xpoly[:, :, 1] = 1
    sh = R.poly_mul(a, xpoly)
    for l, q in enumerate(qs):
        want = np.concatenate([(q - a[l, :, -1:]) % q, a[l, :, :-1]], axis=1)
        assert np.array_equal(sh[l], want)
    if N <= 128:
        for l, q in enumerate(qs):
            assert c[l, 0].tolist() == O.schoolbook_negacyclic(a[l, 0].tolist(), b[l, 0].tolist(), q)


# ---- NTT-domain inner product and automorphism oracles ------------------
@pytest.mark.parametrize("N,bits", [(16, 30), (64, 50), (256, 60)])
def test_automorphism_oracles_agree_with_definition(N, bits):
    q = O.ntt_primes(bits, 1)[0]
    psi = O.find_psi(q, N)
    rng = random.Random(N)
    a = O.random_poly(rng, N, q)
    for k in (3, 5, 2 * N - 1, N + 1, 1):
        c = O.automorphism_coeff(a, k, q)
        # definition: (sigma_k a)(x) = a(x^k) at every root of X^N + 1
        for j in (0, 1, N // 2, N - 1):
            pt = pow(psi, 2 * j + 1, q)
            assert O._horner(c, pt, q) == O._horner(a, pow(pt, k, q), q)
        # the NTT-domain permutation commutes with the transform
        assert O.automorphism_ntt(O.ntt_definition(a, q, psi), k) == O.ntt_definition(c, q, psi)
    # sigma_k is a ring homomorphism: sigma(a*b) = sigma(a)*sigma(b)
    b = O.random_poly(rng, N, q)
    assert O.automorphism_coeff(O.schoolbook_negacyclic(a, b, q), 5, q) == \
        O.schoolbook_negacyclic(O.automorphism_coeff(a, 5, q), O.automorphism_coeff(b, 5, q), q)


def test_mul_acc_oracle_is_the_sum_of_reference_style_products():
    q = O.ntt_primes(60, 1)[0]
    rng = random.Random(7)
    A = [O.random_poly(rng, 32, q) for _ in range(5)]
    B = [O.random_poly(rng, 32, q) for _ in range(5)]
    acc = [0] * 32
    for a, b in zip(A, B):
        acc = O.vec_add(acc, O.vec_mul(a, b, q), q)   # the pinned restatements of arithmetic.py:3-13
    assert O.mul_acc(A, B, q) == acc
