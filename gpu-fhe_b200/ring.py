"""Roots of unity, RNS ring context and the HBM-resident plan cache.

The reference's ``primitive.py`` is the intended home of the primitive roots of unity but holds
only a placeholder (/root/reference/primitive.py:3-4).  The functions here are what the repo-root
``primitive.py`` re-exports.  Root convention: psi = x^((q-1)/2N) mod q for
the smallest integer x >= 2 such that psi^N = -1 (mod q).
"""
from __future__ import annotations

import contextlib
import ctypes
import threading
from typing import Iterable, List, Optional, Sequence, Tuple

from . import _capi

MAX_MODULUS_BITS = 62
MIN_LOG_N, MAX_LOG_N = 4, 20


def is_probable_prime(n: int) -> bool:
    """Deterministic Miller-Rabin for n < 3.3e24."""
    n = int(n)
    if n < 2:
        return False
    bases = (2, 3, 5, 7, 11, 13, 17, 19, 23, 29, 31, 37)
    for p in bases:
        if n % p == 0:
            return n == p
    d, s = n - 1, 0
    while d % 2 == 0:
        d, s = d // 2, s + 1
    for a in bases:
        x = pow(a, d, n)
        if x == 1 or x == n - 1:
            continue
        for _ in range(s - 1):
            x = x * x % n
            if x == n - 1:
                break
        else:
            return False
    return True


def find_psi(q: int, N: int) -> int:
    """Primitive 2N-th root of unity mod q (smallest-base convention)."""
    q, N = int(q), int(N)
    if N < 1 or N & (N - 1):
        raise ValueError("N must be a power of two")
    if (q - 1) % (2 * N):
        raise ValueError("modulus %d is not 1 mod 2N (N=%d)" % (q, N))
    e = (q - 1) // (2 * N)
    for x in range(2, 1 << 16):
        psi = pow(x, e, q)
        if pow(psi, N, q) == q - 1:
            return psi
    raise ValueError("no primitive 2N-th root found for q=%d (is it prime?)" % q)


def ntt_primes(bits: int, count: int, N: int = 1 << 17) -> List[int]:
    """The `count` largest primes below 2**bits with q = 1 mod 2N, descending."""
    step = 2 * int(N)
    q = ((1 << bits) - 1) // step * step + 1
    out: List[int] = []
    while len(out) < count and q > step:
        if is_probable_prime(q):
            out.append(q)
        q -= step
    if len(out) < count:
        raise ValueError("not enough %d-bit primes = 1 mod %d" % (bits, step))
    return out


class Plan:
    """Owner of one gfhe_plan_t (per-limb constants + twiddle tables on one device)."""

    def __init__(self, device: int, N: int, moduli: Sequence[int], psis: Optional[Sequence[int]]):
        self.device, self.N, self.moduli = int(device), int(N), tuple(int(q) for q in moduli)
        self.logN = self.N.bit_length() - 1 if N else 0
        self.L = len(self.moduli)
        self.psis = tuple(int(p) for p in psis) if psis is not None else None
        arr_q = (ctypes.c_uint64 * self.L)(*self.moduli)
        arr_p = (ctypes.c_uint64 * self.L)(*self.psis) if self.psis is not None else None
        h = ctypes.c_void_p()
        _capi.check(_capi.lib.gfhe_plan_create(ctypes.byref(h), self.device, self.logN, self.L, arr_q, arr_p),
                    "gfhe_plan_create")
        self.handle = h

    def close(self):
        if getattr(self, "handle", None):
            _capi.lib.gfhe_plan_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


_plans = {}
_plans_lock = threading.Lock()


def get_plan(device: int, N: int, moduli: Sequence[int]) -> Plan:
    """Cached plan for (device, N, moduli).  N == 0: modulus set only (vec_* ops)."""
    moduli = tuple(int(q) for q in moduli)
    key = (int(device), int(N), moduli)
    with _plans_lock:
        pl = _plans.get(key)
        if pl is None:
            for q in moduli:
                if not 2 <= q < (1 << MAX_MODULUS_BITS):
                    raise ValueError("modulus %d outside [2, 2^%d)" % (q, MAX_MODULUS_BITS))
            psis = None
            if N:
                logn = int(N).bit_length() - 1
                if (1 << logn) != N or not MIN_LOG_N <= logn <= MAX_LOG_N:
                    raise ValueError("N must be a power of two in [2^%d, 2^%d], got %r" % (MIN_LOG_N, MAX_LOG_N, N))
                psis = [find_psi(q, N) for q in moduli]
            pl = Plan(device, N, moduli, psis)
            _plans[key] = pl
        return pl


def clear_plans() -> None:
    with _plans_lock:
        for pl in _plans.values():
            pl.close()
        _plans.clear()


# ---- ring context: how NTT(x) / iNTT(x), which take no modulus argument, learn the moduli -------
_ctx = threading.local()


def _norm_moduli(moduli) -> Tuple[int, ...]:
    if moduli is None:
        raise ValueError("moduli required")
    if hasattr(moduli, "tolist"):
        moduli = moduli.tolist()
    if isinstance(moduli, (int,)) or not isinstance(moduli, Iterable):
        moduli = [moduli]
    flat: List[int] = []

    def rec(v):
        if isinstance(v, (list, tuple)):
            for w in v:
                rec(w)
        else:
            flat.append(int(v))
    rec(list(moduli))
    if not flat:
        raise ValueError("empty modulus list")
    return tuple(flat)


def set_ring(moduli) -> None:
    """Install the RNS modulus chain used by NTT(x) / iNTT(x) when no MOD= is given."""
    _ctx.moduli = None if moduli is None else _norm_moduli(moduli)


def get_ring() -> Optional[Tuple[int, ...]]:
    return getattr(_ctx, "moduli", None)


@contextlib.contextmanager
def ring(moduli):
    """``with ring([q0, q1, ...]): X = NTT(x)``"""
    prev = get_ring()
    set_ring(moduli)
    try:
        yield
    finally:
        _ctx.moduli = prev


def resolve_moduli(MOD) -> Tuple[int, ...]:
    if MOD is not None:
        return _norm_moduli(MOD)
    m = get_ring()
    if m is None:
        raise RuntimeError(
            "NTT/iNTT need the ring moduli: pass MOD=... or call primitive.set_ring([...]) / use "
            "`with primitive.ring([...])` (the reference's identity stub is not reproduced)")
    return m
