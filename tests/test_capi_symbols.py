"""CPU suite, part 2: the C-ABI library loads and exports every symbol include/*.h declares
(no compute calls: there is no GPU here)."""
import ctypes
import glob
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    names = set()
    for h in glob.glob(os.path.join(ROOT, "include", "*.h")):
        text = re.sub(r"/\*.*?\*/", "", open(h).read(), flags=re.S)
        names |= set(re.findall(r"\b(gfhe_\w+)\s*\(", text))
    return names


def test_header_declares_the_expected_surface():
    names = declared_symbols()
    for must in ("gfhe_plan_create", "gfhe_plan_destroy", "gfhe_ntt_forward", "gfhe_ntt_inverse", "gfhe_poly_mul",
                 "gfhe_poly_mul_workspace_bytes", "gfhe_vec_add", "gfhe_vec_sub", "gfhe_vec_mul",
                 "gfhe_ntt_host", "gfhe_poly_mul_host", "gfhe_abi_version", "gfhe_error_string"):
        assert must in names


def test_library_exports_every_declared_symbol(built_lib):
    lib = ctypes.CDLL(built_lib)
    for name in sorted(declared_symbols()):
        assert hasattr(lib, name), "missing export: " + name


def test_binding_covers_header_and_abi_version(built_lib):
    import arithmetic  # noqa: F401  (loads gpu_fhe_b200)
    from gpu_fhe_b200 import _capi
    assert set(_capi.SYMBOLS) == declared_symbols()
    assert _capi.lib.gfhe_abi_version() == _capi.ABI_VERSION == 6
    assert _capi.error_string(0) == "success"
    assert "2N" in _capi.error_string(-4)


def test_argument_errors_need_no_gpu(built_lib):
    import arithmetic  # noqa: F401
    from gpu_fhe_b200 import _capi
    h = ctypes.c_void_p()
    q = (ctypes.c_uint64 * 1)(97)
    psi = (ctypes.c_uint64 * 1)(3)
    # validation happens before any CUDA call
    assert _capi.lib.gfhe_plan_create(None, 0, 4, 1, q, psi) == -1
    assert _capi.lib.gfhe_plan_create(ctypes.byref(h), 0, 3, 1, q, psi) == -2          # log2 N too small
    big = (ctypes.c_uint64 * 1)(1 << 62)
    assert _capi.lib.gfhe_plan_create(ctypes.byref(h), 0, 4, 1, big, psi) == -3        # modulus too wide
    q2 = (ctypes.c_uint64 * 1)(101)
    assert _capi.lib.gfhe_plan_create(ctypes.byref(h), 0, 4, 1, q2, psi) == -4         # 101 != 1 mod 32
    one = (ctypes.c_uint64 * 1)(1)
    assert _capi.lib.gfhe_plan_create(ctypes.byref(h), 0, 4, 1, q, one) == -5          # 1 is not a 32nd root of -1
    assert _capi.lib.gfhe_ntt_forward(None, 0, 1, 1, None, None, None) == -1
    assert _capi.lib.gfhe_poly_mul_workspace_bytes(None, 1, 1) == 0
    # plan images: garbage is refused before any CUDA call
    assert _capi.lib.gfhe_plan_import(ctypes.byref(h), 0, b"not a plan image" * 8, 128) == -14
    assert _capi.lib.gfhe_plan_import(ctypes.byref(h), 0, b"short", 5) == -14
    assert _capi.lib.gfhe_plan_export_bytes(None) == 0
    assert "aligned" in _capi.error_string(-13) and "image" in _capi.error_string(-14)
