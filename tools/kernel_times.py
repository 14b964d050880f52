#!/usr/bin/env python
"""Per-kernel CUDA-event times of the config-4 poly-mul and forward NTT (gfhe_profile_* hooks).  With GFHE_B200_LIB set
it times an experiment build of the library (see tools/exp_variants.sh); results of such builds are not checked."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import polynomial as P
from gpu_fhe_b200 import _capi
N, L, B, bits = 65536, 32, 16, int(os.environ.get("BITS", "60"))
moduli = P.ntt_primes(bits, L); plan = P.make_plan(N, moduli, device=0); lib = _capi.lib
rng = np.random.default_rng(0)
h = np.stack([rng.integers(0, q, (B, N), dtype=np.uint64) for q in moduli]).view(np.int64)
PACKED = os.environ.get("PACKED") == "1"      # packed 32-bit layout (BITS=30 only)
if PACKED: h = h.astype(np.int32)
a = torch.from_numpy(h).cuda(); b = torch.from_numpy(h[::-1].copy()).cuda(); c = torch.empty_like(a)
f_mul, f_fwd, f_inv = (lib.gfhe_poly_mul_u32, lib.gfhe_ntt_forward_u32, lib.gfhe_ntt_inverse_u32) if PACKED else (lib.gfhe_poly_mul, lib.gfhe_ntt_forward, lib.gfhe_ntt_inverse)
wsb = lib.gfhe_poly_mul_workspace_bytes(plan.handle, L, B); ws = torch.empty(wsb, dtype=torch.uint8, device="cuda")
st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
def pm(): f_mul(plan.handle, 0, L, B, a.data_ptr(), b.data_ptr(), c.data_ptr(), ws.data_ptr(), wsb, st)
def nt(): f_fwd(plan.handle, 0, L, B, a.data_ptr(), c.data_ptr(), st)
def it(): f_inv(plan.handle, 0, L, B, a.data_ptr(), c.data_ptr(), st)
for name, f in (("polymul", pm), ("ntt", nt), ("intt", it)):
    for _ in range(5): f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); [f() for _ in range(50)]; e1.record(); torch.cuda.synchronize()
    total = e0.elapsed_time(e1) / 50
    lib.gfhe_profile_enable(plan.handle, 1); [f() for _ in range(20)]
    names = (ctypes.c_char_p * 8)(); tot = (ctypes.c_double * 8)(); cnt = (ctypes.c_uint64 * 8)()
    nk = lib.gfhe_profile_read(plan.handle, 8, names, tot, cnt); lib.gfhe_profile_enable(plan.handle, 0)
    print(os.environ.get("TAG", "base"), "bits=%d%s pipe=%s" % (bits, " packed" if PACKED else "", os.environ.get("GFHE_PIPE", "-")), name, "step %.4f ms |" % total, "  ".join("%s %.4f" % (names[i].decode(), tot[i] / cnt[i]) for i in range(nk)), flush=True)
