"""gfhe_poly_mul_static -- the product entry point that allocates nothing (its scratch is a file-scope device array of
GFHE_STATIC_WS_BYTES inside the library; include/gpufhe_b200.h).  It has no counterpart in the reference: it was added
for callers that cannot call cudaMalloc.  Called through the C-ABI and compared,
bit for bit (the tolerance of every integer test here), with the existing gfhe_poly_mul on the same inputs and with the
oracle, on two calls in a row that share the scratch.  Collected last on purpose."""
import ctypes

import numpy as np
import pytest
import torch

from oracle import cref
from oracle import oracle as O

pytestmark = pytest.mark.gpu

ESTATICWS, ENOALLOC = -15, -16


def dev(a, device):
    return torch.from_numpy(np.ascontiguousarray(a).view(np.int64)).to(device)


def host(t):
    return t.cpu().numpy().view(np.uint64)


def rand_lbn(seed, moduli, B, N):
    rng = np.random.default_rng(seed)
    a = np.stack([rng.integers(0, q, (B, N), dtype=np.uint64) for q in moduli])
    a[:, :, :3] = np.array(moduli, dtype=np.uint64).reshape(len(moduli), 1, 1) - 1      # extremes
    a[:, :, 3] = 0
    return a


def stream_of(device):
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def existing(lib, plan, L, B, ta, tb, device):
    """gfhe_poly_mul with a caller-allocated workspace: the entry point the new one is checked against"""
    tc = torch.zeros_like(ta)
    need = lib.gfhe_poly_mul_workspace_bytes(plan.handle, L, B)
    ws = torch.empty(max(need, 16), dtype=torch.uint8, device=device)
    assert lib.gfhe_poly_mul(plan.handle, 0, L, B, ta.data_ptr(), tb.data_ptr(), tc.data_ptr(), ws.data_ptr(), need, stream_of(device)) == 0
    torch.cuda.synchronize()
    return tc


def static(lib, plan, L, B, ta, tb, device):
    tc = torch.zeros_like(ta)
    rc = lib.gfhe_poly_mul_static(plan.handle, 0, L, B, ta.data_ptr(), tb.data_ptr(), tc.data_ptr(), stream_of(device))
    torch.cuda.synchronize()
    return rc, tc


# 60-bit (Lazy60), 61-bit (Harvey), 50-bit (Fp50) and 30-bit (Shoup32) limbs; N on both stage splits, N = 4096 needs no scratch
@pytest.mark.parametrize("bits,logn,L,B", [(60, 16, 2, 3), (61, 13, 3, 2), (50, 14, 2, 5), (30, 16, 3, 4), (60, 12, 2, 3), (60, 17, 1, 2)])
def test_static_scratch_product_matches_existing_entry_point_twice_in_a_row(cuda_device, bits, logn, L, B):
    import polynomial as P
    from gpu_fhe_b200 import _capi
    lib = _capi.lib
    N = 1 << logn
    moduli = P.ntt_primes(bits, L, N=max(N, 1 << 17))
    plan = P.make_plan(N, moduli, device=cuda_device.index or 0)
    R = cref.Ring(N, moduli, [O.find_psi(q, N) for q in moduli])
    # two different jobs, back to back on the same scratch: the second finds the first one's words in it
    jobs = [(rand_lbn(7 * logn + bits, moduli, B, N), rand_lbn(7 * logn + bits + 1, moduli, B, N))]
    jobs.append((jobs[0][1][:, ::-1].copy(), rand_lbn(7 * logn + bits + 2, moduli, B, N)))
    for a, b in jobs:
        ta, tb = dev(a, cuda_device), dev(b, cuda_device)
        want = existing(lib, plan, L, B, ta, tb, cuda_device)
        rc, got = static(lib, plan, L, B, ta, tb, cuda_device)
        assert rc == 0, _capi.error_string(rc)
        assert torch.equal(got, want)
        assert np.array_equal(host(got), R.poly_mul(a, b, fast=True, nthreads=4))
        assert torch.equal(ta, dev(a, cuda_device)) and torch.equal(tb, dev(b, cuda_device))      # inputs untouched


def test_static_scratch_limit_and_no_allocation(cuda_device):
    """[4,16,65536] is exactly GFHE_STATIC_WS_BYTES and runs; one polynomial more is refused with GFHE_ESTATICWS before
    anything is launched; repeated calls leave the device's free memory where it was (nothing is allocated or freed)."""
    import polynomial as P
    from gpu_fhe_b200 import _capi
    lib = _capi.lib
    N, L, B = 65536, 4, 16
    assert lib.gfhe_static_workspace_bytes() == L * B * N * 8 == 32 << 20
    moduli = P.ntt_primes(60, L)
    plan = P.make_plan(N, moduli, device=cuda_device.index or 0)
    a, b = rand_lbn(1, moduli, B + 1, N), rand_lbn(2, moduli, B + 1, N)
    ta, tb = dev(a[:, :B], cuda_device), dev(b[:, :B], cuda_device)
    want = existing(lib, plan, L, B, ta, tb, cuda_device)
    rc, got = static(lib, plan, L, B, ta, tb, cuda_device)          # first call on this device resolves the array's address
    assert rc == 0 and torch.equal(got, want)
    tc = torch.zeros_like(ta)
    free0 = torch.cuda.mem_get_info(cuda_device)[0]
    for _ in range(3):
        assert lib.gfhe_poly_mul_static(plan.handle, 0, L, B, ta.data_ptr(), tb.data_ptr(), tc.data_ptr(), stream_of(cuda_device)) == 0
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info(cuda_device)[0] >= free0
    assert torch.equal(tc, want)
    # a limb range of the plan: limbs 1..2, compared with the same range of the full result
    tc2 = torch.zeros_like(ta[1:3])
    assert lib.gfhe_poly_mul_static(plan.handle, 1, 2, B, ta[1:3].data_ptr(), tb[1:3].data_ptr(), tc2.data_ptr(), stream_of(cuda_device)) == 0
    torch.cuda.synchronize()
    assert torch.equal(tc2, want[1:3])
    # beyond the limit: refused, output untouched
    tb1, ta1 = dev(b, cuda_device), dev(a, cuda_device)
    rc, got = static(lib, plan, L, B + 1, ta1, tb1, cuda_device)
    assert rc == ESTATICWS and "GFHE_STATIC_WS_BYTES" in _capi.error_string(rc)
    assert not got.any()
    # argument checks shared with gfhe_poly_mul
    assert lib.gfhe_poly_mul_static(plan.handle, 0, L, B, ta.data_ptr(), tb.data_ptr(), ta.data_ptr(), stream_of(cuda_device)) == -1
    assert lib.gfhe_poly_mul_static(plan.handle, 3, 2, B, ta.data_ptr(), tb.data_ptr(), tc.data_ptr(), stream_of(cuda_device)) == -7
    assert lib.gfhe_poly_mul_static(plan.handle, 0, L, B, ta.data_ptr() + 8, tb.data_ptr(), tc.data_ptr(), stream_of(cuda_device)) == -13


def test_static_scratch_refuses_pipelined_plans(cuda_device, monkeypatch):
    import polynomial as P
    from gpu_fhe_b200 import _capi
    monkeypatch.setenv("GFHE_PIPE", "1")
    P.clear_plans()
    try:
        moduli = P.ntt_primes(60, 1)
        plan = P.make_plan(8192, moduli, device=cuda_device.index or 0)
        a = dev(rand_lbn(3, moduli, 1, 8192), cuda_device)
        rc, got = static(_capi.lib, plan, 1, 1, a, a.clone(), cuda_device)
        assert rc == ENOALLOC and not got.any()
    finally:
        P.clear_plans()
