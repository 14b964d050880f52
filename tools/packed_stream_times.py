#!/usr/bin/env python
"""CUDA-event times of the packed-layout streaming kernels at config-4 shape (32 limbs x 16 x 65536, 30-bit moduli)."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import polynomial as P
from gpu_fhe_b200 import _capi
N, L, B = 65536, 32, 16
moduli = P.ntt_primes(30, L); plan = P.make_plan(N, moduli, device=0); lib = _capi.lib
red = torch.tensor(moduli, dtype=torch.int64, device="cuda").view(L, 1, 1)
g = torch.Generator(device="cuda"); g.manual_seed(1)
a = (torch.randint(0, 1 << 30, (4, L, B, N), device="cuda", generator=g) % red).to(torch.int32)
b = (torch.randint(0, 1 << 30, (4, L, B, N), device="cuda", generator=g) % red).to(torch.int32)
c = torch.empty((L, B, N), dtype=torch.int32, device="cuda")
st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
inner = B * N
cases = {
    "vec_add_i32 (12 B/res)": (lambda: lib.gfhe_vec_add_i32(plan.handle, 0, L, inner, a[0].data_ptr(), b[0].data_ptr(), c.data_ptr(), st), 12),
    "vec_mul_i32 (12 B/res)": (lambda: lib.gfhe_vec_mul_i32(plan.handle, 0, L, inner, a[1].data_ptr(), b[1].data_ptr(), c.data_ptr(), st), 12),
    "mul_acc_u32 x1 (12 B/res)": (lambda: lib.gfhe_ntt_mul_acc_u32(plan.handle, 0, L, inner, 1, a.data_ptr(), b.data_ptr(), L * inner, c.data_ptr(), st), 12),
    "mul_acc_u32 x4 (36 B/res)": (lambda: lib.gfhe_ntt_mul_acc_u32(plan.handle, 0, L, inner, 4, a.data_ptr(), b.data_ptr(), L * inner, c.data_ptr(), st), 36),
    "automorphism_u32 ntt (8 B/res)": (lambda: lib.gfhe_automorphism_u32(plan.handle, 0, L, B, 5, 1, a[2].data_ptr(), c.data_ptr(), st), 8),
    "automorphism_u32 coeff (8 B/res)": (lambda: lib.gfhe_automorphism_u32(plan.handle, 0, L, B, 5, 0, a[3].data_ptr(), c.data_ptr(), st), 8),
}
for name, (f, bpr) in cases.items():
    for _ in range(5):
        assert f() == 0
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); [f() for _ in range(30)]; e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 30
    print("%-34s %.4f ms  %.0f GB/s" % (name, ms, bpr * L * inner / ms / 1e6), flush=True)
