"""Marshalling between the reference-level Python signatures and the C ABI.

Device buffers are PyTorch CUDA tensors (int64 or uint64, contiguous); torch is used for
allocation, streams and copies only -- every residue is computed by the CUDA kernels behind
include/gpufhe_b200.h.  Host inputs (numpy arrays, CPU tensors) are copied to the GPU, never
computed on the CPU: there is no CPU fallback, and without CUDA every call raises.
"""
from __future__ import annotations

import ctypes
from typing import Optional, Sequence, Tuple

import numpy as np
import torch

from . import _capi, ring

_VEC_FN = {0: "gfhe_vec_add", 1: "gfhe_vec_sub", 2: "gfhe_vec_mul"}
_INT_DTYPES = (torch.int64, torch.uint64)
_PACKED_DTYPES = (torch.int32, torch.uint32)   # packed layout: CUDA tensors of 32-bit words, moduli below 2^30 only


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
        if isinstance(x, torch.Tensor):
            if x.dtype in _PACKED_DTYPES:
                self.packed = True
            elif x.dtype not in _INT_DTYPES:
                raise TypeError("expected an int64/uint64 (or packed int32/uint32) tensor of residues, got %s" % x.dtype)
            if x.is_cuda:
                self.t = x.contiguous()
            else:
                _require_cuda()
                self.kind = "cpu_tensor"
                self.t = x.contiguous().to(device or torch.device("cuda", torch.cuda.current_device()), non_blocking=False)
        else:
            _require_cuda()
            a = np.asarray(x)
            self.kind = "numpy"
            self.np_dtype = a.dtype
            if a.dtype == object:
                a = np.array([int(v) & 0xFFFFFFFFFFFFFFFF for v in a.ravel()], dtype=np.uint64).reshape(a.shape)
            elif a.dtype not in (np.int64, np.uint64):
                if not np.issubdtype(a.dtype, np.integer):
                    raise TypeError("expected integer residues, got dtype %s" % a.dtype)
                a = a.astype(np.int64)
            a = np.ascontiguousarray(a)
            self.host = a
            self.t = torch.from_numpy(a.view(np.int64)).to(device or torch.device("cuda", torch.cuda.current_device()))
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

def _vec_moduli(MOD, shape: Tuple[int, ...]) -> Tuple[Tuple[int, ...], int]:
    """Interpret MOD with numpy broadcasting rules against `shape`.

    Supported: a scalar, or an array that (after left-padding its shape with 1s) is 1 everywhere
    except possibly along axis 0 -- i.e. one modulus per leading index, e.g. shape [L,1,1] for
    [L,B,N] operands.  Returns (moduli, elements per modulus).
    """
    numel = 1
    for s in shape:
        numel *= s
    if isinstance(MOD, torch.Tensor):
        MOD = MOD.detach().cpu().numpy()
    m = np.asarray(MOD, dtype=object) if not isinstance(MOD, np.ndarray) else MOD
    if m.ndim == 0 or m.size == 1:
        return (int(m.reshape(-1)[0]),), numel
    if m.ndim > len(shape):
        raise ValueError("MOD of shape %s does not broadcast to operands of shape %s" % (m.shape, shape))
    padded = (1,) * (len(shape) - m.ndim) + tuple(m.shape)
    if padded[0] != shape[0] or any(d != 1 for d in padded[1:]):
        raise ValueError(
            "MOD of shape %s: only a scalar or one modulus per leading index (shape %s) is supported"
            % (tuple(m.shape), (shape[0],) + (1,) * (len(shape) - 1)))
    return tuple(int(v) for v in m.reshape(-1)), numel // shape[0]


def vec_op(op: int, a, b, MOD):
    A = _Arg(a)
    B = _Arg(b, A.t.device)
    _same_layout(A, B)
    if B.t.device != A.t.device:
        raise ValueError("operands live on different devices")
    moduli, inner = _vec_moduli(MOD, A.shape)
    out = torch.empty_like(A.t)
    if A.t.numel() == 0:
        return A.wrap(out)
    dev = A.t.device
    plan = ring.get_plan(dev.index, 0, moduli)
    name = _VEC_FN[op] + ("_i32" if A.packed else "")
    _capi.check(getattr(_capi.lib, name)(plan.handle, 0, len(moduli), inner, _ptr(A.t), _ptr(B.t), _ptr(out), _stream(dev)), name)
    return A.wrap(out)


# ---- transforms (/root/reference/arithmetic.py:15-19) and product -----------------------------------

def _same_layout(a: "_Arg", b: "_Arg") -> None:
    if a.packed != b.packed:
        raise TypeError("operands must use the same word size (both 64-bit or both packed 32-bit)")


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
    _require_cuda()
    moduli = ring._norm_moduli(moduli)
    L, B, N = _lbn(a.shape, len(moduli))
    plan = ring.get_plan(device, N, moduli)
    _host_u64(a), _host_u64(b), _host_u64(out)
    _capi.check(_capi.lib.gfhe_poly_mul_host(plan.handle, B, a.ctypes.data, b.ctypes.data, out.ctypes.data),
                "gfhe_poly_mul_host")
    return out


def ntt_host(x: np.ndarray, out: np.ndarray, moduli, inverse: bool = False, device: int = 0) -> np.ndarray:
    _require_cuda()
    moduli = ring._norm_moduli(moduli)
    L, B, N = _lbn(x.shape, len(moduli))
    plan = ring.get_plan(device, N, moduli)
    _host_u64(x), _host_u64(out)
    _capi.check(_capi.lib.gfhe_ntt_host(plan.handle, int(inverse), B, x.ctypes.data, out.ctypes.data), "gfhe_ntt_host")
    return out
