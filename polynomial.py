"""Drop-in for /root/reference/" polynomial.py" (the reference file name has a leading space and
cannot be imported by name; this one can).

    poly_add(a, b, MOD)   " polynomial.py":3-5 -- component-wise sum of two ciphertext-like pairs.
                          The reference computes both sums and drops them (returns None); this
                          returns them, which is signature-compatible.
    poly_sub(a, b, MOD)   same shape of call, difference
    poly_mul(a, b, MOD)   new: negacyclic product in Z_q[X]/(X^N+1) per RNS limb,
                          iNTT(NTT(a) * NTT(b)) as one fused GPU pipeline
    poly_mul_pre(a, B)    a * b with B = NTT(b) cached by the caller (b's forward transform skipped)
    poly_mul_ntt(A, B)    NTT-domain pointwise product of operands that are already transformed
    poly_mac_ntt(As, Bs)  NTT-domain inner product sum_t As[t] * Bs[t] (key-switching style), one reduction per residue
    poly_mul_pair(a, b)   product of two 2-component pairs (a0 + a1 s)(b0 + b1 s) -> 3 components, 4 forward + 3 inverse
                          transforms with the cross term as one NTT-domain inner product
    automorphism(x, k)    Galois map a(X) -> a(X^k), coefficient domain or (ntt=True) NTT domain
"""
from primitive import *  # noqa: F401,F403
from arithmetic import _d

__all__ = ["np", "vec_add", "vec_sub", "vec_mul", "NTT", "iNTT", "XXX",
           "find_psi", "ntt_primes", "is_probable_prime", "set_ring", "get_ring", "ring", "make_plan", "clear_plans",
           "save_plan", "load_plan",
           "poly_add", "poly_sub", "poly_mul", "poly_mul_pre", "poly_mul_ntt", "poly_mac_ntt", "automorphism", "poly_mul_pair"]


def poly_add(a, b, MOD):
    return (vec_add(a[0], b[0], MOD), vec_add(a[1], b[1], MOD))   # noqa: F405


def poly_sub(a, b, MOD):
    return (vec_sub(a[0], b[0], MOD), vec_sub(a[1], b[1], MOD))   # noqa: F405


def poly_mul(a, b, MOD=None):
    """a, b: [L, B, N] (or [B, N] / [N] with a single modulus); MOD: scalar, list of L moduli, an
    [L,1,1] array, or None to use the installed ring.  Coefficients must be canonical residues in [0, q)
    (see arithmetic.py: reduce centred / negative coefficients with vec_add first)."""
    assert a.shape == b.shape
    return _d.poly_mul(a, b, MOD)


def poly_mul_pre(a, b_ntt, MOD=None):
    """poly_mul(a, b) given b_ntt = NTT(b): the fused pipeline without b's forward transform."""
    assert a.shape == b_ntt.shape
    return _d.poly_mul_pre(a, b_ntt, MOD)


def poly_mul_ntt(a, b, MOD=None):
    """NTT(a*b) from NTT(a), NTT(b): [L, B, N] canonical residues in, same out."""
    assert a.shape == b.shape
    return _d.mul_acc(a, b, MOD, False)


def poly_mac_ntt(a, b, MOD=None):
    """sum_t a[t] * b[t] in the NTT domain.  a, b: [K, L, B, N] stacks or length-K sequences of [L, B, N]."""
    return _d.mul_acc(a, b, MOD, True)


def automorphism(x, k, *, ntt=False, MOD=None):
    """a(X) -> a(X^k) for odd k; with ntt=True x and the result are in NTT(x)'s evaluation order, and
    automorphism(NTT(a), k, ntt=True) == NTT(automorphism(a, k))."""
    return _d.automorphism(x, k, MOD, ntt)


def poly_mul_pair(a, b, MOD=None):
    """(a[0] + a[1] s) * (b[0] + b[1] s) = c0 + c1 s + c2 s^2 over Z_q[X]/(X^N+1): returns (c0, c1, c2) with
    c0 = a0 b0, c1 = a0 b1 + a1 b0, c2 = a1 b1.  Same pair indexing as poly_add (" polynomial.py":3-5)."""
    assert a[0].shape == a[1].shape == b[0].shape == b[1].shape
    A0, A1, B0, B1 = (NTT(x, MOD=MOD) for x in (a[0], a[1], b[0], b[1]))   # noqa: F405
    c0 = iNTT(_d.mul_acc(A0, B0, MOD, False), MOD=MOD)                       # noqa: F405
    c1 = iNTT(_d.mul_acc([A0, A1], [B1, B0], MOD, True), MOD=MOD)            # noqa: F405
    c2 = iNTT(_d.mul_acc(A1, B1, MOD, False), MOD=MOD)                       # noqa: F405
    return (c0, c1, c2)
