"""Marshalling between the reference-level Python signatures and the C ABI.

Device buffers are PyTorch CUDA tensors (int64 or uint64, contiguous); torch is used for
allocation, streams and copies only -- every residue is computed by the CUDA kernels behind
include/gpufhe_b200.h.  Host inputs (numpy arrays, CPU tensors) are copied to the GPU, never
computed on the CPU: there is no CPU fallback, and without CUDA every call raises.
"""
from __future__ import annotations

import ctypes
import os
from typing import Optional, Sequence, Tuple

import numpy as np
import torch

from . import _capi, ring

_VEC_FN = {0: "gfhe_vec_add", 1: "gfhe_vec_sub", 2: "gfhe_vec_mul"}
_INT_DTYPES = (torch.int64, torch.uint64)
_PACKED_DTYPES = (torch.int32, torch.uint32)   # packed layout: CUDA tensors of 32-bit words, moduli below 2^30 only
_CHECK_RANGE = os.environ.get("GFHE_CHECK_RANGE", "0") == "1"


def _require_cuda() -> None:
    if not torch.cuda.is_available():
        raise RuntimeError("gpufhe_b200: no CUDA device available and there is no CPU fallback")


def _stream(device: torch.device) -> ctypes.c_void_p:
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def _ptr(t: torch.Tensor) -> ctypes.c_void_p:
    return ctypes.c_void_p(t.data_ptr())


class _Arg:
    """One operand normalised to a contiguous CUDA int64/uint64 tensor + how to hand results back."""

    def __init__(self, x, device: Optional[torch.device] = None):
        self.kind = "cuda"
        self.np_dtype = None
        self.packed = False
        self.unsigned = False     # 64-bit words are unsigned values (uint64 containers): >= 2^63 is not negative
        if isinstance(x, torch.Tensor):
            if x.dtype in _PACKED_DTYPES:
                self.packed = True
            elif x.dtype not in _INT_DTYPES:
                raise TypeError("expected an int64/uint64 (or packed int32/uint32) tensor of residues, got %s" % x.dtype)
            self.unsigned = x.dtype == torch.uint64
            if x.is_cuda:
                self.t = x.contiguous()
            else:
                _require_cuda()
                self.kind = "cpu_tensor"
                self.t = x.contiguous().to(device or torch.device("cuda", torch.cuda.current_device()), non_blocking=False)
        else:
            a = np.asarray(x)
            self.kind = "numpy"
            self.np_dtype = a.dtype
            if a.dtype == object:
                # exact Python ints: they must fit a 64-bit word, signed or unsigned -- silently truncating (or
                # mis-signing) them would break the exactness the reference has in this mode
                flat = [int(v) for v in a.ravel()]
                lo, hi = (min(flat), max(flat)) if flat else (0, 0)
                if lo >= -(1 << 63) and hi < (1 << 63):
                    a = np.array(flat, dtype=np.int64).reshape(a.shape)
                elif lo >= 0 and hi < (1 << 64):
                    a = np.array(flat, dtype=np.uint64).reshape(a.shape)
                else:
                    raise ValueError("object-dtype residues must fit one 64-bit word: all in [-2^63, 2^63) or all in "
                                     "[0, 2^64); got values in [%d, %d]" % (lo, hi))
            elif a.dtype not in (np.int64, np.uint64):
                if not np.issubdtype(a.dtype, np.integer):
                    raise TypeError("expected integer residues, got dtype %s" % a.dtype)
                a = a.astype(np.int64)
            _require_cuda()
            a = np.ascontiguousarray(a)
            self.host = a
            self.unsigned = a.dtype == np.uint64
            self.t = torch.from_numpy(a.view(np.int64)).to(device or torch.device("cuda", torch.cuda.current_device()))
        if self.t.data_ptr() % 16:
            # a contiguous view at an odd offset (x.contiguous() returns it as it is): the transform kernels use 16-byte
            # accesses and the C ABI refuses such pointers (GFHE_EALIGN), so give them an aligned copy
            self.t = self.t.clone()
        self.shape = tuple(self.t.shape)

    def wrap(self, out: torch.Tensor):
        """Return `out` (a CUDA tensor) in the container type of the original operand."""
        if self.kind == "cuda":
            return out
        if self.kind == "cpu_tensor":
            return out.cpu()
        h = out.cpu().numpy()
        if self.np_dtype == object:
            return h.view(np.uint64).astype(object)
        if self.np_dtype == np.uint64:
            return h.view(np.uint64)
        return h.view(np.int64).astype(self.np_dtype, copy=False)


# ---- element-wise ops (/root/reference/arithmetic.py:3-13) ----------------------------------------

def _vec_moduli(MOD, shape: Tuple[int, ...]):
    """Interpret MOD with numpy's broadcasting rules against operands of `shape` (/root/reference/arithmetic.py:5,9,13
    accept anything `%` broadcasts).

    Returns (moduli, out_shape, perm, inner): the kernels take [len(moduli), inner] rows with one modulus per row, so the
    axes along which MOD varies are moved to the front (`perm`, None when they already lead -- e.g. MOD of shape [L,1,1]
    against [L,B,N]) and flattened; `out_shape` is the broadcast result shape (larger than `shape` when MOD has more
    axes or the operands have size-1 axes where MOD does not).
    """
    if isinstance(MOD, torch.Tensor):
        MOD = MOD.detach().cpu().numpy()
    m = np.asarray(MOD, dtype=object) if not isinstance(MOD, np.ndarray) else MOD
    if m.ndim == 0 or m.size == 1:
        numel = 1
        for d in shape:
            numel *= d
        return (int(m.reshape(-1)[0]),), tuple(shape), None, numel
    out_shape = tuple(int(d) for d in np.broadcast_shapes(tuple(shape), m.shape))   # ValueError like numpy's when they do not broadcast
    nd = len(out_shape)
    mshape = (1,) * (nd - m.ndim) + tuple(m.shape)
    varying = [d for d in range(nd) if mshape[d] != 1]
    rest = [d for d in range(nd) if mshape[d] == 1]
    inner = 1
    for d in rest:
        inner *= out_shape[d]
    perm = varying + rest
    return tuple(int(v) for v in m.reshape(-1)), out_shape, (None if perm == list(range(nd)) else perm), inner


def vec_op(op: int, a, b, MOD):
    A = _Arg(a)
    B = _Arg(b, A.t.device)
    _same_layout(A, B)
    if B.t.device != A.t.device:
        raise ValueError("operands live on different devices")
    unsigned = A.unsigned or B.unsigned
    if A.unsigned != B.unsigned and not A.packed:
        # one uint64 and one int64 container: fine as long as the signed one holds no negative value (the bit patterns
        # then mean the same thing either way); a negative value next to values >= 2^63 has no common 64-bit reading
        signed = B.t if A.unsigned else A.t
        if bool((signed.view(torch.int64) < 0).any()):
            raise ValueError("cannot mix a uint64 operand with an int64 operand that holds negative values")
    moduli, out_shape, perm, inner = _vec_moduli(MOD, A.shape)
    # torch implements few operators for the unsigned dtypes: move the words around as their signed twins
    twin = {torch.uint64: torch.int64, torch.uint32: torch.int32}
    ta, tb = (t.view(twin[t.dtype]) if t.dtype in twin else t for t in (A.t, B.t))
    if out_shape != A.shape:                      # MOD has more axes than the operands, or they have size-1 axes it fills
        ta, tb = ta.expand(out_shape), tb.expand(out_shape)
    if perm is not None:                          # MOD varies along non-leading axes: bring them to the front
        ta, tb = ta.permute(perm), tb.permute(perm)
    ta, tb = ta.contiguous(), tb.contiguous()
    out = torch.empty_like(ta)
    if ta.numel():
        dev = ta.device
        plan = ring.get_plan(dev.index, 0, moduli)
        name = _VEC_FN[op] + ("_i32" if A.packed else ("_u64" if unsigned else ""))
        _capi.check(getattr(_capi.lib, name)(plan.handle, 0, len(moduli), inner, _ptr(ta), _ptr(tb), _ptr(out), _stream(dev)), name)
    if perm is not None:
        inv = [0] * len(perm)
        for k, d in enumerate(perm):
            inv[d] = k
        out = out.permute(inv).contiguous()
    return A.wrap(out.view(A.t.dtype))


# ---- transforms (/root/reference/arithmetic.py:15-19) and product -----------------------------------

def _same_layout(a: "_Arg", b: "_Arg") -> None:
    if a.packed != b.packed:
        raise TypeError("operands must use the same word size (both 64-bit or both packed 32-bit)")


def _check_canonical(t: torch.Tensor, moduli: Sequence[int], L: int, what: str) -> None:
    """Opt-in (GFHE_CHECK_RANGE=1) test of the [0, q) precondition of the transform / product entry points.  Unlike
    vec_*, which accept any integer, NTT / iNTT / poly_mul / poly_mul_pre / mul_acc / automorphism take canonical
    residues: the lazy-reduction range schedule of the kernels starts from q (centred or negative coefficients must go
    through vec_add(x, 0, MOD) first).  One reduction per call plus a sync, so it is a debugging aid, not the default."""
    if not _CHECK_RANGE or t.numel() == 0:
        return
    v = t.view(torch.int64) if t.dtype == torch.uint64 else (t.view(torch.int32) if t.dtype == torch.uint32 else t)
    q = torch.tensor([int(m) for m in moduli], dtype=torch.int64, device=t.device)
    rows = v.reshape(L, -1) if L > 1 else v.reshape(1, -1)
    bad = (rows < 0) | (rows >= q.view(-1, 1))
    if bool(bad.any()):
        raise ValueError("%s: residues outside [0, q) (the transform entry points need canonical residues)" % what)


def _lbn(shape: Sequence[int], L: int) -> Tuple[int, int, int]:
    """Map an operand shape (..., N) onto the canonical [L, B, N]."""
    if len(shape) == 0:
        raise ValueError("operand must have at least one dimension (the N coefficients)")
    N = int(shape[-1])
    lead = 1
    for s in shape[:-1]:
        lead *= int(s)
    if L == 1:
        return 1, lead, N
    if len(shape) < 2 or int(shape[0]) != L:
        raise ValueError("with %d moduli the leading dimension must be %d (layout [L, B, N]); got shape %s"
                         % (L, L, tuple(shape)))
    return L, lead // L, N


def ntt(x, MOD, inverse: bool):
    moduli = ring.resolve_moduli(MOD)
    X = _Arg(x)
    L, B, N = _lbn(X.shape, len(moduli))
    dev = X.t.device
    plan = ring.get_plan(dev.index, N, moduli)
    out = torch.empty_like(X.t)
    if B == 0:
        return X.wrap(out)
    _check_canonical(X.t, moduli, L, "iNTT" if inverse else "NTT")
    if X.packed:
        fn = _capi.lib.gfhe_ntt_inverse_u32 if inverse else _capi.lib.gfhe_ntt_forward_u32
    else:
        fn = _capi.lib.gfhe_ntt_inverse if inverse else _capi.lib.gfhe_ntt_forward
    _capi.check(fn(plan.handle, 0, L, B, _ptr(X.t), _ptr(out), _stream(dev)), "gfhe_ntt")
    return X.wrap(out)


def poly_mul(a, b, MOD):
    moduli = ring.resolve_moduli(MOD)
    A = _Arg(a)
    Bv = _Arg(b, A.t.device)
    if A.shape != Bv.shape:
        raise AssertionError("operands must have equal shapes")   # same error class as arithmetic.py:4
    L, B, N = _lbn(A.shape, len(moduli))
    dev = A.t.device
    plan = ring.get_plan(dev.index, N, moduli)
    out = torch.empty_like(A.t)
    if B == 0:
        return A.wrap(out)
    _same_layout(A, Bv)
    _check_canonical(A.t, moduli, L, "poly_mul"), _check_canonical(Bv.t, moduli, L, "poly_mul")
    ws_bytes = _capi.lib.gfhe_poly_mul_workspace_bytes(plan.handle, L, B) // (2 if A.packed else 1)
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev) if ws_bytes else None
    fn = _capi.lib.gfhe_poly_mul_u32 if A.packed else _capi.lib.gfhe_poly_mul
    _capi.check(fn(plan.handle, 0, L, B, _ptr(A.t), _ptr(Bv.t), _ptr(out),
                                        _ptr(ws) if ws is not None else None, ws_bytes, _stream(dev)), "gfhe_poly_mul")
    return A.wrap(out)


def poly_mul_pre(a, b_ntt, MOD):
    """iNTT(NTT(a) * b_ntt): a in the coefficient domain, b_ntt = NTT(b) cached by the caller."""
    moduli = ring.resolve_moduli(MOD)
    A = _Arg(a)
    Bv = _Arg(b_ntt, A.t.device)
    if A.shape != Bv.shape:
        raise AssertionError("operands must have equal shapes")
    L, B, N = _lbn(A.shape, len(moduli))
    dev = A.t.device
    plan = ring.get_plan(dev.index, N, moduli)
    out = torch.empty_like(A.t)
    if B == 0:
        return A.wrap(out)
    _same_layout(A, Bv)
    fn = _capi.lib.gfhe_poly_mul_pre_u32 if A.packed else _capi.lib.gfhe_poly_mul_pre
    _capi.check(fn(plan.handle, 0, L, B, _ptr(A.t), _ptr(Bv.t), _ptr(out), _stream(dev)), "gfhe_poly_mul_pre")
    return A.wrap(out)


# ---- NTT-domain-resident ops ------------------------------------

def mul_acc(a, b, MOD, stacked: bool):
    """sum_t a[t] * b[t] mod q per limb.  a, b: [K, L, B, N] stacks (or sequences of K [L, B, N] operands) of canonical
    NTT-domain residues; returns [L, B, N].  A [L, B, N] pair (no leading K) is the plain pointwise product."""
    moduli = ring.resolve_moduli(MOD)
    if isinstance(a, (list, tuple)):
        a = torch.stack([_Arg(x).t for x in a]) if not isinstance(a[0], np.ndarray) else np.stack(a)
    if isinstance(b, (list, tuple)):
        b = torch.stack([_Arg(x).t for x in b]) if not isinstance(b[0], np.ndarray) else np.stack(b)
    A = _Arg(a)
    Bv = _Arg(b, A.t.device)
    _same_layout(A, Bv)
    if A.shape != Bv.shape:
        raise AssertionError("operands must have equal shapes")
    L = len(moduli)
    shape = A.shape
    if stacked:
        if len(shape) < 2 or shape[0] == 0:
            raise ValueError("expected a non-empty leading term axis ([K, L, B, N]); got shape %s" % (shape,))
        K, item = int(shape[0]), shape[1:]
    else:
        K, item = 1, shape
    Lq, Bq, N = _lbn(item, L)
    inner = Bq * N
    dev = A.t.device
    plan = ring.get_plan(dev.index, 0, moduli)
    out = torch.empty(item, dtype=A.t.dtype, device=dev)
    if inner == 0:
        return A.wrap(out)
    fn = _capi.lib.gfhe_ntt_mul_acc_u32 if A.packed else _capi.lib.gfhe_ntt_mul_acc
    _capi.check(fn(plan.handle, 0, Lq, inner, K, _ptr(A.t), _ptr(Bv.t), Lq * inner, _ptr(out), _stream(dev)), "gfhe_ntt_mul_acc")
    return A.wrap(out)


def automorphism(x, k: int, MOD, ntt_domain: bool):
    """a(X) -> a(X^k), k odd, on every limb-polynomial of x ([L, B, N]); out of place."""
    moduli = ring.resolve_moduli(MOD)
    X = _Arg(x)
    L, B, N = _lbn(X.shape, len(moduli))
    k = int(k) % (2 * N)
    if k % 2 == 0:
        raise ValueError("Galois element k must be odd, got %d" % k)
    dev = X.t.device
    plan = ring.get_plan(dev.index, N, moduli)
    out = torch.empty_like(X.t)
    if B == 0:
        return X.wrap(out)
    fn = _capi.lib.gfhe_automorphism_u32 if X.packed else _capi.lib.gfhe_automorphism
    _capi.check(fn(plan.handle, 0, L, B, k, int(bool(ntt_domain)), _ptr(X.t), _ptr(out), _stream(dev)), "gfhe_automorphism")
    return X.wrap(out)


# ---- host-buffer entry points (what bench.py's e2e leg times) ---------------------------------------

def _host_u64(a: np.ndarray) -> np.ndarray:
    if a.dtype not in (np.uint64, np.int64) or not a.flags.c_contiguous:
        raise TypeError("host entry points take C-contiguous int64/uint64 arrays")
    return a


def poly_mul_host(a: np.ndarray, b: np.ndarray, out: np.ndarray, moduli, device: int = 0) -> np.ndarray:
    """c = a * b with [L,B,N] HOST arrays; H2D / kernels / D2H are pipelined inside the C call."""
    moduli = ring._norm_moduli(moduli)
    _host_u64(a), _host_u64(b), _host_u64(out)
    if not (a.shape == b.shape == out.shape):     # the C side copies L*B*N words from / to each of the three
        raise AssertionError("a, b and out must have equal shapes, got %s, %s, %s" % (a.shape, b.shape, out.shape))
    if not out.flags.writeable:
        raise ValueError("out must be writable")
    L, B, N = _lbn(a.shape, len(moduli))
    if L * B * N != a.size:
        raise ValueError("shape %s does not hold %d limbs" % (a.shape, len(moduli)))
    _require_cuda()
    plan = ring.get_plan(device, N, moduli)
    _capi.check(_capi.lib.gfhe_poly_mul_host(plan.handle, B, a.ctypes.data, b.ctypes.data, out.ctypes.data),
                "gfhe_poly_mul_host")
    return out


def host_stats(moduli, N: int, device: int = 0) -> dict:
    """Where the last *_host call on this ring spent its time (gfhe_host_stats): stream-busy milliseconds of the H2D
    copies, the kernels and the D2H copies, the bytes each way, and the host staging time of pageable callers."""
    plan = ring.get_plan(device, N, ring._norm_moduli(moduli))
    buf = (ctypes.c_double * 10)()
    n = _capi.lib.gfhe_host_stats(plan.handle, buf, 10)
    keys = ("h2d_ms", "kernel_ms", "d2h_ms", "h2d_bytes", "d2h_bytes", "wall_ms", "stage_in_ms", "stage_out_ms", "staged", "groups")
    return {k: buf[i] for i, k in enumerate(keys[:n])}


def ntt_host(x: np.ndarray, out: np.ndarray, moduli, inverse: bool = False, device: int = 0) -> np.ndarray:
    moduli = ring._norm_moduli(moduli)
    _host_u64(x), _host_u64(out)
    if x.shape != out.shape:
        raise AssertionError("x and out must have equal shapes, got %s, %s" % (x.shape, out.shape))
    if not out.flags.writeable:
        raise ValueError("out must be writable")
    L, B, N = _lbn(x.shape, len(moduli))
    if L * B * N != x.size:
        raise ValueError("shape %s does not hold %d limbs" % (x.shape, len(moduli)))
    _require_cuda()
    plan = ring.get_plan(device, N, moduli)
    _capi.check(_capi.lib.gfhe_ntt_host(plan.handle, int(inverse), B, x.ctypes.data, out.ctypes.data), "gfhe_ntt_host")
    return out
