"""Roots of unity, RNS ring context and the HBM-resident plan cache.

The reference's ``primitive.py`` is the intended home of the primitive roots of unity but holds
only a placeholder (/root/reference/primitive.py:3-4).  The functions here are what the repo-root
``primitive.py`` re-exports.  Root convention: psi = x^((q-1)/2N) mod q for
the smallest integer x >= 2 such that psi^N = -1 (mod q).
"""
from __future__ import annotations

import collections
import contextlib
import ctypes
import hashlib
import os
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

    def __init__(self, device: int, N: int, moduli: Sequence[int], psis: Optional[Sequence[int]], _handle=None):
        self.device, self.N, self.moduli = int(device), int(N), tuple(int(q) for q in moduli)
        self.logN = self.N.bit_length() - 1 if N else 0
        self.L = len(self.moduli)
        self.psis = tuple(int(p) for p in psis) if psis is not None else None
        if _handle is not None:          # adopted from gfhe_plan_import (load_plan)
            self.handle = _handle
            return
        arr_q = (ctypes.c_uint64 * self.L)(*self.moduli)
        arr_p = (ctypes.c_uint64 * self.L)(*self.psis) if self.psis is not None else None
        h = ctypes.c_void_p()
        _capi.check(_capi.lib.gfhe_plan_create(ctypes.byref(h), self.device, self.logN, self.L, arr_q, arr_p),
                    "gfhe_plan_create")
        self.handle = h

    def to_bytes(self) -> bytes:
        """The plan image (gfhe_plan_export): header, moduli, roots, per-limb constants and the twiddle tables as they sit
        in HBM, checksummed.  load_plan / Plan.from_bytes turn it back into a resident plan without any root search or
        table generation."""
        n = _capi.lib.gfhe_plan_export_bytes(self.handle)
        buf = ctypes.create_string_buffer(n)
        _capi.check(_capi.lib.gfhe_plan_export(self.handle, buf, n), "gfhe_plan_export")
        return buf.raw

    @classmethod
    def from_bytes(cls, image: bytes, device: int) -> "Plan":
        h = ctypes.c_void_p()
        _capi.check(_capi.lib.gfhe_plan_import(ctypes.byref(h), int(device), image, len(image)), "gfhe_plan_import")
        dev, logn, L = ctypes.c_int(), ctypes.c_uint32(), ctypes.c_uint32()
        _capi.check(_capi.lib.gfhe_plan_info(h, ctypes.byref(dev), ctypes.byref(logn), ctypes.byref(L)), "gfhe_plan_info")
        q = (ctypes.c_uint64 * L.value)()
        p = (ctypes.c_uint64 * L.value)()
        _capi.lib.gfhe_plan_moduli(h, q, p, L.value)
        N = (1 << logn.value) if logn.value else 0
        return cls(dev.value, N, list(q), list(p) if N else None, _handle=h)

    def close(self):
        if getattr(self, "handle", None):
            _capi.lib.gfhe_plan_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def save_plan(plan: Plan, path: str) -> str:
    """Write the plan image to `path` (atomically: temp file + rename)."""
    tmp = "%s.tmp.%d" % (path, os.getpid())
    with open(tmp, "wb") as f:
        f.write(plan.to_bytes())
    os.replace(tmp, path)
    return path


def load_plan(path: str, device: int = 0) -> Plan:
    """Make the plan saved at `path` resident on `device`.  Raises GpuFheError (GFHE_EFORMAT) for a file that is not a
    plan image of this library, is truncated, or fails its checksum."""
    with open(path, "rb") as f:
        return Plan.from_bytes(f.read(), device)


def _cache_path(N: int, moduli: Tuple[int, ...]) -> Optional[str]:
    """File of the on-disk table cache for (N, moduli), or None when GFHE_PLAN_CACHE_DIR is not set."""
    root = os.environ.get("GFHE_PLAN_CACHE_DIR")
    if not root or not N:
        return None
    key = hashlib.sha256(("abi%d|N%d|" % (_capi.ABI_VERSION, N) + ",".join(str(q) for q in moduli)).encode()).hexdigest()[:32]
    return os.path.join(root, "plan_%s.gfhe" % key)


_plans = {}
_vec_plans: "collections.OrderedDict" = collections.OrderedDict()   # N == 0 plans (vec_* modulus sets): LRU, see get_plan
_VEC_PLAN_CAP = 64
_plans_lock = threading.Lock()


def get_plan(device: int, N: int, moduli: Sequence[int]) -> Plan:
    """Cached plan for (device, N, moduli).  N == 0: modulus set only (vec_* ops).

    NTT plans stay until clear_plans().  Modulus-set plans are cheap but one is made per distinct MOD a caller ever
    passes to vec_*, so that cache is an LRU of _VEC_PLAN_CAP entries (an evicted plan is destroyed when its last user
    drops it).  With GFHE_PLAN_CACHE_DIR set, NTT plans are also saved to / loaded from disk (save_plan / load_plan)."""
    moduli = tuple(int(q) for q in moduli)
    key = (int(device), int(N), moduli)
    with _plans_lock:
        if not N:
            pl = _vec_plans.get(key)
            if pl is not None:
                _vec_plans.move_to_end(key)
                return pl
        else:
            pl = _plans.get(key)
            if pl is not None:
                return pl
        for q in moduli:
            if not 2 <= q < (1 << MAX_MODULUS_BITS):
                raise ValueError("modulus %d outside [2, 2^%d)" % (q, MAX_MODULUS_BITS))
        if not N:
            pl = Plan(device, 0, moduli, None)
            _vec_plans[key] = pl
            while len(_vec_plans) > _VEC_PLAN_CAP:
                _vec_plans.popitem(last=False)      # not closed here: another thread may be mid-call on it; __del__ frees it
            return pl
        logn = int(N).bit_length() - 1
        if (1 << logn) != N or not MIN_LOG_N <= logn <= MAX_LOG_N:
            raise ValueError("N must be a power of two in [2^%d, 2^%d], got %r" % (MIN_LOG_N, MAX_LOG_N, N))
        path = _cache_path(N, moduli)
        pl = None
        if path and os.path.exists(path):
            try:
                pl = load_plan(path, device)
                if pl.N != N or pl.moduli != moduli:
                    pl.close()
                    pl = None
            except _capi.GpuFheError:
                pl = None                            # stale or damaged image: rebuild and overwrite
        if pl is None:
            pl = Plan(device, N, moduli, [find_psi(q, N) for q in moduli])
            if path:
                try:
                    os.makedirs(os.path.dirname(path), exist_ok=True)
                    save_plan(pl, path)
                except OSError:
                    pass
        _plans[key] = pl
        return pl


def clear_plans() -> None:
    with _plans_lock:
        for pl in list(_plans.values()) + list(_vec_plans.values()):
            pl.close()
        _plans.clear()
        _vec_plans.clear()


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
