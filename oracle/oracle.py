"""CPU oracle for the negacyclic-NTT / RNS poly-mul hot path.  TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline /
``--impl reference`` legs may import this module.  The product path
(``arithmetic.py`` / ``primitive.py`` / ``polynomial.py`` -> C-ABI -> CUDA) never does.

Parity status
-------------
* ``vec_add`` / ``vec_sub`` / ``vec_mul`` restate ``/root/reference/arithmetic.py:3-13``
  (``(a+b)%MOD``, ``(a-b)%MOD``, ``(a*b)%MOD`` after a shape assert) in exact Python-int
  arithmetic, i.e. the reference run on ``dtype=object`` arrays -- the only mode in which
  the reference is mathematically exact for >31-bit moduli.  PINNED: tests/golden/vec_ops.npz
  was produced by importing the reference itself (tests/golden/make_golden.py).
* ``NTT`` / ``iNTT`` / ``poly_mul``: **parity unpinned** by the reference.
  ``/root/reference/arithmetic.py:15-19`` are identity stubs, the reference has no
  roots of unity (``primitive.py:3-4`` prints "XXX") and no poly-mul
  (``" polynomial.py":3-5``), no tests and no golden vectors.  These functions are
  therefore written from the mathematical definition:

      NTT(x)[k]  = x(psi^(2*brv(k)+1)) mod q         (bit-reversed evaluation order)
      iNTT       = its inverse, output fully reduced
      poly_mul   = product in Z_q[X]/(X^N+1)

  and cross-checked three independent ways (definition-level evaluation, schoolbook
  negacyclic convolution, Kronecker substitution), none of which shares code or tables
  with the CUDA path.

Everything here uses Python ints (arbitrary precision), so there is no overflow mode.
"""
from __future__ import annotations

import random
from typing import List, Sequence

# --------------------------------------------------------------------------------------
# element-wise ops -- restate /root/reference/arithmetic.py:3-13
# --------------------------------------------------------------------------------------

def _shape(x):
    return getattr(x, "shape", (len(x),))


def vec_add(a, b, MOD):
    """arithmetic.py:3-5 -- assert a.shape == b.shape; return (a + b) % MOD."""
    assert _shape(a) == _shape(b)
    return [(int(x) + int(y)) % int(MOD) for x, y in zip(a, b)]


def vec_sub(a, b, MOD):
    """arithmetic.py:7-9 -- (a - b) % MOD, Python-style non-negative remainder."""
    assert _shape(a) == _shape(b)
    return [(int(x) - int(y)) % int(MOD) for x, y in zip(a, b)]


def vec_mul(a, b, MOD):
    """arithmetic.py:11-13 -- (a * b) % MOD, exact (object-dtype semantics)."""
    assert _shape(a) == _shape(b)
    return [(int(x) * int(y)) % int(MOD) for x, y in zip(a, b)]


# --------------------------------------------------------------------------------------
# roots of unity -- the convention SURVEY.md section 8c pins
# --------------------------------------------------------------------------------------

def is_probable_prime(n: int) -> bool:
    if n < 2:
        return False
    small = (2, 3, 5, 7, 11, 13, 17, 19, 23, 29, 31, 37)
    for p in small:
        if n % p == 0:
            return n == p
    d, s = n - 1, 0
    while d % 2 == 0:
        d //= 2
        s += 1
    for a in small:  # deterministic for n < 3.3e24
        x = pow(a, d, n)
        if x in (1, n - 1):
            continue
        for _ in range(s - 1):
            x = x * x % n
            if x == n - 1:
                break
        else:
            return False
    return True


def find_psi(q: int, N: int) -> int:
    """psi = pow(x, (q-1)//(2N), q) for the smallest integer x >= 2 with psi^N == -1 (mod q)."""
    assert N >= 1 and N & (N - 1) == 0
    assert (q - 1) % (2 * N) == 0, "q must be 1 mod 2N"
    e = (q - 1) // (2 * N)
    x = 2
    while True:
        psi = pow(x, e, q)
        if pow(psi, N, q) == q - 1:
            return psi
        x += 1


def ntt_primes(bits: int, count: int, log2_step: int = 18) -> List[int]:
    """Largest primes below 2**bits with q == 1 (mod 2**log2_step), descending."""
    step = 1 << log2_step
    q = ((1 << bits) - 1) // step * step + 1
    out = []
    while len(out) < count:
        if q < (1 << bits) and is_probable_prime(q):
            out.append(q)
        q -= step
    return out


def brv(k: int, bits: int) -> int:
    r = 0
    for _ in range(bits):
        r = (r << 1) | (k & 1)
        k >>= 1
    return r


# --------------------------------------------------------------------------------------
# definition-level transforms (O(N^2); N <= ~2^10)
# --------------------------------------------------------------------------------------

def _horner(x: Sequence[int], pt: int, q: int) -> int:
    acc = 0
    for c in reversed(x):
        acc = (acc * pt + c) % q
    return acc


def ntt_definition(x: Sequence[int], q: int, psi: int) -> List[int]:
    """NTT(x)[k] = x(psi^(2*brv(k)+1)) mod q, straight from the definition."""
    N = len(x)
    bits = N.bit_length() - 1
    x = [int(v) % q for v in x]
    return [_horner(x, pow(psi, 2 * brv(k, bits) + 1, q), q) for k in range(N)]


def intt_definition(X: Sequence[int], q: int, psi: int) -> List[int]:
    """x[j] = N^-1 * sum_k X[k] * psi^(-(2*brv(k)+1)*j): the inverse of ntt_definition."""
    N = len(X)
    bits = N.bit_length() - 1
    ninv = pow(N, -1, q)
    pinv = pow(psi, -1, q)
    pts = [pow(pinv, 2 * brv(k, bits) + 1, q) for k in range(N)]
    out = []
    for j in range(N):
        acc = 0
        for k in range(N):
            acc += int(X[k]) * pow(pts[k], j, q)
        out.append(acc % q * ninv % q)
    return out


def schoolbook_negacyclic(a: Sequence[int], b: Sequence[int], q: int) -> List[int]:
    """c[k] = sum_{i+j=k} a_i b_j - sum_{i+j=k+N} a_i b_j (mod q): the canonical product."""
    N = len(a)
    assert len(b) == N
    c = [0] * N
    a = [int(v) for v in a]
    b = [int(v) for v in b]
    for i, ai in enumerate(a):
        if ai == 0:
            continue
        for j, bj in enumerate(b):
            k = i + j
            if k < N:
                c[k] += ai * bj
            else:
                c[k - N] -= ai * bj
    return [v % q for v in c]


def schoolbook_coeff(a: Sequence[int], b: Sequence[int], q: int, k: int) -> int:
    """One output coefficient of the negacyclic product, O(N)."""
    N = len(a)
    acc = 0
    for i in range(N):
        j = k - i
        if j >= 0:
            acc += int(a[i]) * int(b[j])
        else:
            acc -= int(a[i]) * int(b[j + N])
    return acc % q


def kronecker_negacyclic(a: Sequence[int], b: Sequence[int], q: int) -> List[int]:
    """Negacyclic product by Kronecker substitution: needs no roots, no NTT.

    Pack each operand into one big int with slots wide enough for any product
    coefficient (< N * q^2), multiply once, unpack 2N slots, fold c[k] - c[k+N], reduce.
    """
    N = len(a)
    slot = 2 * q.bit_length() + N.bit_length() + 1
    slot = (slot + 7) // 8 * 8
    nb = slot // 8
    A = int.from_bytes(b"".join(int(v).to_bytes(nb, "little") for v in a), "little")
    Bv = int.from_bytes(b"".join(int(v).to_bytes(nb, "little") for v in b), "little")
    P = (A * Bv).to_bytes(2 * N * nb, "little")
    full = [int.from_bytes(P[i * nb:(i + 1) * nb], "little") for i in range(2 * N)]
    return [(full[k] - full[k + N]) % q for k in range(N)]


# --------------------------------------------------------------------------------------
# O(N log N) transforms, written as the recursive CRT tree  Z_q[X]/(X^{2t} - r^2) ->
# Z_q[X]/(X^t - r) x Z_q[X]/(X^t + r)  (no precomputed tables; shares nothing with the
# CUDA path's bit-reversed twiddle tables).
# --------------------------------------------------------------------------------------

def ntt_fast(x: Sequence[int], q: int, psi: int) -> List[int]:
    N = len(x)
    a = [int(v) % q for v in x]

    def rec(seg: List[int], e: int) -> List[int]:
        # seg is an element of Z_q[X]/(X^n - psi^e); leaves come out in the order that
        # makes leaf k the evaluation at psi^(2*brv(k)+1).
        n = len(seg)
        if n == 1:
            return seg
        t = n // 2
        # X^n - psi^e = (X^t - psi^(e/2)) (X^t + psi^(e/2)),  -psi^(e/2) = psi^(e/2 + N)
        r = pow(psi, e // 2, q)
        lo = [(seg[j] + r * seg[j + t]) % q for j in range(t)]
        hi = [(seg[j] - r * seg[j + t]) % q for j in range(t)]
        return rec(lo, e // 2) + rec(hi, e // 2 + N)

    # X^N + 1 = X^N - psi^N
    return rec(a, N)


def intt_fast(X: Sequence[int], q: int, psi: int) -> List[int]:
    N = len(X)
    inv2 = pow(2, -1, q)

    def rec(vals: List[int], e: int) -> List[int]:
        n = len(vals)
        if n == 1:
            return [int(vals[0]) % q]
        t = n // 2
        lo = rec(vals[:t], e // 2)
        hi = rec(vals[t:], e // 2 + N)
        rinv = pow(psi, -(e // 2), q)
        first = [(lo[j] + hi[j]) * inv2 % q for j in range(t)]
        second = [(lo[j] - hi[j]) * inv2 % q * rinv % q for j in range(t)]
        return first + second

    return rec(list(X), N)


def poly_mul(a: Sequence[int], b: Sequence[int], q: int, psi: int | None = None) -> List[int]:
    """Negacyclic product via the transforms above: iNTT(NTT(a) * NTT(b))."""
    if psi is None:
        psi = find_psi(q, len(a))
    A = ntt_fast(a, q, psi)
    Bv = ntt_fast(b, q, psi)
    return intt_fast([x * y % q for x, y in zip(A, Bv)], q, psi)


# --------------------------------------------------------------------------------------
# cheap spot checks usable at full size (always used as a pair)
# --------------------------------------------------------------------------------------

def check_point_eval(a, b, c, q: int, psi: int, k: int) -> bool:
    """c(psi_k) == a(psi_k) * b(psi_k) at the root psi_k = psi^(2k+1) of X^N + 1."""
    pt = pow(psi, 2 * k + 1, q)
    return _horner(c, pt, q) == _horner(a, pt, q) * _horner(b, pt, q) % q


def ntt_point(x, q: int, psi: int, k: int) -> int:
    """The value NTT(x)[k] must have: x(psi^(2*brv(k)+1))."""
    bits = len(x).bit_length() - 1
    return _horner([int(v) % q for v in x], pow(psi, 2 * brv(k, bits) + 1, q), q)


def random_poly(rng: random.Random, N: int, q: int) -> List[int]:
    return [rng.randrange(q) for _ in range(N)]


# ---- NTT-domain inner product and Galois automorphism -------------------
# The reference has neither; these restate the definitions in Python ints.

def mul_acc(a_terms: Sequence[Sequence[int]], b_terms: Sequence[Sequence[int]], q: int) -> List[int]:
    """out[i] = sum_t a_terms[t][i] * b_terms[t][i] mod q."""
    n = len(a_terms[0])
    return [sum(int(a[i]) * int(b[i]) for a, b in zip(a_terms, b_terms)) % q for i in range(n)]


def automorphism_coeff(a: Sequence[int], k: int, q: int) -> List[int]:
    """Coefficients of a(X^k) mod (X^N + 1, q), k odd: a_i X^(ik), with X^N = -1."""
    N = len(a)
    out = [0] * N
    for i, v in enumerate(a):
        e = (i * k) % (2 * N)
        if e < N:
            out[e] = int(v) % q
        else:
            out[e - N] = (-int(v)) % q
    return out


def automorphism_ntt(X: Sequence[int], k: int) -> List[int]:
    """The same map on NTT(a) in bit-reversed evaluation order: position j holds a(psi^(2 brv(j) + 1)), and
    (sigma_k a)(psi^e) = a(psi^(k e))."""
    N = len(X)
    bits = N.bit_length() - 1
    out = [0] * N
    for j in range(N):
        e = (k * (2 * brv(j, bits) + 1)) % (2 * N)
        out[j] = X[brv((e - 1) // 2, bits)]
    return out
