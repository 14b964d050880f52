"""ctypes binding of the C ABI declared in include/gpufhe_b200.h.

There is NO fallback: if ``libgpufhe_b200.so`` has not been built
(``python -c "import __graft_entry__ as g; g.build()"`` or ``make -C gpu-fhe_b200/csrc``)
importing this module raises, and so does every reference-level function.
"""
from __future__ import annotations

import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("GFHE_B200_LIB") or os.path.join(_HERE, "libgpufhe_b200.so")   # override: timing experiments only

u64p = ctypes.POINTER(ctypes.c_uint64)
i64p = ctypes.POINTER(ctypes.c_int64)
vp = ctypes.c_void_p
u32 = ctypes.c_uint32

# name -> (restype, argtypes); must list every symbol include/gpufhe_b200.h declares
SYMBOLS = {
    "gfhe_abi_version": (ctypes.c_int, []),
    "gfhe_error_string": (ctypes.c_char_p, [ctypes.c_int]),
    "gfhe_plan_create": (ctypes.c_int, [ctypes.POINTER(vp), ctypes.c_int, u32, u32, u64p, u64p]),
    "gfhe_plan_destroy": (ctypes.c_int, [vp]),
    "gfhe_plan_info": (ctypes.c_int, [vp, ctypes.POINTER(ctypes.c_int), ctypes.POINTER(u32), ctypes.POINTER(u32)]),
    "gfhe_plan_moduli": (ctypes.c_int, [vp, u64p, u64p, u32]),
    "gfhe_plan_export_bytes": (ctypes.c_size_t, [vp]),
    "gfhe_plan_export": (ctypes.c_int, [vp, vp, ctypes.c_size_t]),
    "gfhe_plan_import": (ctypes.c_int, [ctypes.POINTER(vp), ctypes.c_int, vp, ctypes.c_size_t]),
    "gfhe_host_stats": (ctypes.c_int, [vp, ctypes.POINTER(ctypes.c_double), u32]),
    "gfhe_ntt_forward": (ctypes.c_int, [vp, u32, u32, u32, vp, vp, vp]),
    "gfhe_ntt_inverse": (ctypes.c_int, [vp, u32, u32, u32, vp, vp, vp]),
    "gfhe_poly_mul_workspace_bytes": (ctypes.c_size_t, [vp, u32, u32]),
    "gfhe_poly_mul": (ctypes.c_int, [vp, u32, u32, u32, vp, vp, vp, vp, ctypes.c_size_t, vp]),
    "gfhe_vec_add": (ctypes.c_int, [vp, u32, u32, ctypes.c_uint64, vp, vp, vp, vp]),
    "gfhe_vec_sub": (ctypes.c_int, [vp, u32, u32, ctypes.c_uint64, vp, vp, vp, vp]),
    "gfhe_vec_mul": (ctypes.c_int, [vp, u32, u32, ctypes.c_uint64, vp, vp, vp, vp]),
    "gfhe_vec_add_u64": (ctypes.c_int, [vp, u32, u32, ctypes.c_uint64, vp, vp, vp, vp]),
    "gfhe_vec_sub_u64": (ctypes.c_int, [vp, u32, u32, ctypes.c_uint64, vp, vp, vp, vp]),
    "gfhe_vec_mul_u64": (ctypes.c_int, [vp, u32, u32, ctypes.c_uint64, vp, vp, vp, vp]),
    "gfhe_poly_mul_pre": (ctypes.c_int, [vp, u32, u32, u32, vp, vp, vp, vp]),
    "gfhe_poly_mul_static": (ctypes.c_int, [vp, u32, u32, u32, vp, vp, vp, vp]),
    "gfhe_static_workspace_bytes": (ctypes.c_size_t, []),
    "gfhe_ntt_forward_u32": (ctypes.c_int, [vp, u32, u32, u32, vp, vp, vp]),
    "gfhe_ntt_inverse_u32": (ctypes.c_int, [vp, u32, u32, u32, vp, vp, vp]),
    "gfhe_poly_mul_u32": (ctypes.c_int, [vp, u32, u32, u32, vp, vp, vp, vp, ctypes.c_size_t, vp]),
    "gfhe_poly_mul_pre_u32": (ctypes.c_int, [vp, u32, u32, u32, vp, vp, vp, vp]),
    "gfhe_vec_add_i32": (ctypes.c_int, [vp, u32, u32, ctypes.c_uint64, vp, vp, vp, vp]),
    "gfhe_vec_sub_i32": (ctypes.c_int, [vp, u32, u32, ctypes.c_uint64, vp, vp, vp, vp]),
    "gfhe_vec_mul_i32": (ctypes.c_int, [vp, u32, u32, ctypes.c_uint64, vp, vp, vp, vp]),
    "gfhe_ntt_mul_acc_u32": (ctypes.c_int, [vp, u32, u32, ctypes.c_uint64, u32, vp, vp, ctypes.c_uint64, vp, vp]),
    "gfhe_automorphism_u32": (ctypes.c_int, [vp, u32, u32, u32, u32, ctypes.c_int, vp, vp, vp]),
    "gfhe_ntt_mul_acc": (ctypes.c_int, [vp, u32, u32, ctypes.c_uint64, u32, vp, vp, ctypes.c_uint64, vp, vp]),
    "gfhe_automorphism": (ctypes.c_int, [vp, u32, u32, u32, u32, ctypes.c_int, vp, vp, vp]),
    "gfhe_peer_alloc": (ctypes.c_int, [ctypes.c_int, ctypes.c_size_t, ctypes.POINTER(vp), ctypes.c_char_p]),
    "gfhe_peer_open": (ctypes.c_int, [ctypes.c_int, ctypes.c_char_p, ctypes.POINTER(vp)]),
    "gfhe_peer_close": (ctypes.c_int, [ctypes.c_int, vp]),
    "gfhe_peer_free": (ctypes.c_int, [ctypes.c_int, vp]),
    "gfhe_push_gather": (ctypes.c_int, [ctypes.c_int, vp, ctypes.c_size_t, ctypes.POINTER(vp), u32, ctypes.c_size_t, u32, vp]),
    "gfhe_ntt_host": (ctypes.c_int, [vp, ctypes.c_int, u32, vp, vp]),
    "gfhe_poly_mul_host": (ctypes.c_int, [vp, u32, vp, vp, vp]),
    "gfhe_debug_pipe_schedule": (ctypes.c_int, [u32, ctypes.POINTER(u32), u32, u32, ctypes.POINTER(u32), u32, ctypes.POINTER(u32)]),
    "gfhe_profile_enable": (ctypes.c_int, [vp, ctypes.c_int]),
    "gfhe_profile_read": (ctypes.c_int, [vp, u32, ctypes.POINTER(ctypes.c_char_p), ctypes.POINTER(ctypes.c_double),
                                         ctypes.POINTER(ctypes.c_uint64)]),
}


class GpuFheError(RuntimeError):
    def __init__(self, code: int, what: str):
        super().__init__("gpufhe_b200: %s failed: %s (code %d)" % (what, error_string(code), code))
        self.code = code


if not os.path.exists(LIB_PATH):
    raise ImportError(
        "gpufhe_b200: %s is missing -- build the CUDA extension first (python -c 'import __graft_entry__ as g; "
        "g.build()'); there is no CPU fallback" % LIB_PATH)

lib = ctypes.CDLL(LIB_PATH)
for _name, (_res, _args) in SYMBOLS.items():
    _fn = getattr(lib, _name)   # AttributeError here = header and library out of sync
    _fn.restype = _res
    _fn.argtypes = _args

ABI_VERSION = 6
if lib.gfhe_abi_version() != ABI_VERSION:
    raise ImportError("gpufhe_b200: ABI version mismatch (library %d, binding %d)" % (lib.gfhe_abi_version(), ABI_VERSION))


def error_string(code: int) -> str:
    return lib.gfhe_error_string(int(code)).decode()


def check(code: int, what: str) -> None:
    if code != 0:
        raise GpuFheError(code, what)
