"""GPU parity of the pipelined single-launch kernels (gpu-fhe_b200/csrc/ntt_pipe.inl; GFHE_PIPE=1 at plan creation):
the same words as the one-launch-per-pass path and as the oracle, for every arithmetic policy, both layouts, sizes on
both stage splits (8 and 12 row stages), and polynomial counts below, around and far above the pipeline lag."""
import numpy as np
import pytest
import torch

from oracle import cref
from oracle import oracle as O

pytestmark = pytest.mark.gpu


def dev(a, device, packed=False):
    h = np.ascontiguousarray(a).view(np.int64)
    if packed:
        h = h.astype(np.int32)
    return torch.from_numpy(h).to(device)


def host(t):
    h = t.cpu().numpy()
    return h.astype(np.uint64) if h.dtype == np.int32 else h.view(np.uint64)


def rand_lbn(seed, moduli, B, N):
    rng = np.random.default_rng(seed)
    a = np.stack([rng.integers(0, q, (B, N), dtype=np.uint64) for q in moduli])
    a[:, :, :3] = np.array(moduli, dtype=np.uint64).reshape(len(moduli), 1, 1) - 1      # extremes
    a[:, :, 3] = 0
    return a


@pytest.fixture
def pipe_on(monkeypatch):
    import primitive as P
    monkeypatch.setenv("GFHE_PIPE", "1")
    P.clear_plans()                      # the knob is read when a plan is created
    yield
    P.clear_plans()


# (bits, packed, extra env): Lazy60, Harvey, Fp50, Shoup32 on the 64-bit layout, Shoup32P packed
POLICIES = [(60, False, {}), (61, False, {}), (50, False, {}), (30, False, {}), (30, True, {}), (60, False, {"GFHE_FORCE_HARVEY": "1"})]


@pytest.mark.parametrize("bits,packed,env", POLICIES)
@pytest.mark.parametrize("logn,L,B", [(13, 2, 3), (14, 3, 2), (16, 2, 2), (17, 1, 3), (18, 1, 1), (13, 4, 150), (16, 5, 13)])
def test_pipelined_kernels_match_oracle(cuda_device, pipe_on, monkeypatch, bits, packed, env, logn, L, B):
    import polynomial as P
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    N = 1 << logn
    moduli = P.ntt_primes(bits, L, N=max(N, 1 << 17))       # q = 1 mod 2N
    a, b = rand_lbn(logn + bits, moduli, B, N), rand_lbn(logn + bits + 1, moduli, B, N)
    R = cref.Ring(N, moduli, [O.find_psi(q, N) for q in moduli])
    ta, tb = dev(a, cuda_device, packed), dev(b, cuda_device, packed)
    with P.ring(moduli):
        X = P.NTT(ta)
        want_X = R.ntt(a, fast=True, nthreads=8)
        assert np.array_equal(host(X), want_X)
        assert torch.equal(P.iNTT(X), ta)
        c = P.poly_mul(ta, tb)
        want_c = R.poly_mul(a, b, fast=True, nthreads=8)
        assert np.array_equal(host(c), want_c)
        assert np.array_equal(host(P.poly_mul_pre(ta, P.NTT(tb))), want_c)
        # twice in a row on the same stream: the synchronisation words must come back zeroed
        assert np.array_equal(host(P.poly_mul(ta, tb)), want_c)
        assert np.array_equal(host(P.NTT(ta)), want_X)


def test_pipelined_in_place_and_streams(cuda_device, pipe_on):
    """in == out through the C ABI (the row pass then reads addresses the column pass's inputs came from), a limb
    sub-range, and two streams at once on one plan (one set of synchronisation words per stream)."""
    import ctypes

    import primitive as P
    from gpu_fhe_b200 import _capi
    N, L, B = 1 << 16, 4, 6
    moduli = P.ntt_primes(60, L)
    plan = P.make_plan(N, moduli, device=cuda_device.index)
    x = rand_lbn(1, moduli, B, N)
    R = cref.Ring(N, moduli, [O.find_psi(q, N) for q in moduli])
    want = R.ntt(x, fast=True, nthreads=8)
    lib = _capi.lib
    t = dev(x, cuda_device)
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    assert lib.gfhe_ntt_forward(plan.handle, 0, L, B, t.data_ptr(), t.data_ptr(), st) == 0
    assert np.array_equal(host(t), want)
    assert lib.gfhe_ntt_inverse(plan.handle, 0, L, B, t.data_ptr(), t.data_ptr(), st) == 0
    assert np.array_equal(host(t), x)
    sub = dev(x[1:3], cuda_device)
    out = torch.empty_like(sub)
    assert lib.gfhe_ntt_forward(plan.handle, 1, 2, B, sub.data_ptr(), out.data_ptr(), st) == 0
    assert np.array_equal(host(out), want[1:3])
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    o1, o2 = torch.empty_like(t), torch.empty_like(t)
    torch.cuda.synchronize()
    for _ in range(3):
        assert lib.gfhe_ntt_forward(plan.handle, 0, L, B, t.data_ptr(), o1.data_ptr(), ctypes.c_void_p(s1.cuda_stream)) == 0
        assert lib.gfhe_ntt_forward(plan.handle, 0, L, B, t.data_ptr(), o2.data_ptr(), ctypes.c_void_p(s2.cuda_stream)) == 0
    torch.cuda.synchronize()
    assert np.array_equal(host(o1), want) and np.array_equal(host(o2), want)


def test_pipelined_config4_full_size(cuda_device, pipe_on):
    import polynomial as P
    N, L, B = 1 << 16, 32, 16
    moduli = P.ntt_primes(60, L)
    a, b = rand_lbn(41, moduli, B, N), rand_lbn(42, moduli, B, N)
    R = cref.Ring(N, moduli, [O.find_psi(q, N) for q in moduli])
    c = P.poly_mul(dev(a, cuda_device), dev(b, cuda_device), list(moduli))
    assert np.array_equal(host(c), R.poly_mul(a, b, fast=True, nthreads=16))


@pytest.mark.parametrize("bits,packed", [(60, False), (50, False), (30, False), (30, True)])
@pytest.mark.parametrize("logn,L,B", [(13, 2, 5), (14, 3, 4), (16, 5, 3), (16, 4, 16), (14, 1, 7), (17, 1, 16)])
def test_two_stream_split_matches_oracle(cuda_device, monkeypatch, bits, packed, logn, L, B):
    """plan.cu split_run: the limb range cut in two halves on two streams (forced here; by default only jobs of 512..4096 tiles
    take it, which the last shape is).  Odd limb counts, every policy, and back-to-back calls that reuse the fork/join events."""
    import polynomial as P
    if not ((L == 4 and B == 16) or (L == 1 and B == 16)):      # those two take the split by size (1024 / 512 tiles)
        monkeypatch.setenv("GFHE_SPLIT", "1")
    P.clear_plans()
    N = 1 << logn
    moduli = P.ntt_primes(bits, L)
    a, b = rand_lbn(3 * logn + bits, moduli, B, N), rand_lbn(3 * logn + bits + 1, moduli, B, N)
    R = cref.Ring(N, moduli, [O.find_psi(q, N) for q in moduli])
    ta, tb = dev(a, cuda_device, packed), dev(b, cuda_device, packed)
    want_c, want_X = R.poly_mul(a, b, fast=True, nthreads=8), R.ntt(a, fast=True, nthreads=8)
    with P.ring(moduli):
        for _ in range(3):
            X = P.NTT(ta)
            c = P.poly_mul(ta, tb)
            back = P.iNTT(X)
            pre = P.poly_mul_pre(ta, P.NTT(tb))
        assert np.array_equal(host(X), want_X)
        assert np.array_equal(host(c), want_c)
        assert torch.equal(back, ta)
        assert np.array_equal(host(pre), want_c)
    P.clear_plans()


@pytest.mark.parametrize("packed", [False, True])
def test_chunk_schedule_of_large_32bit_jobs(cuda_device, packed):
    """plan.cu split_run, large jobs on 32-bit words (> 4096 tiles, >= 4 limbs): chunks of ~2048 tiles alternate between two
    streams.  [6, 48, 65536] is 4608 tiles -> three 2-limb chunks; every word against the oracle, twice in a row."""
    import polynomial as P
    P.clear_plans()
    N, L, B = 1 << 16, 6, 48
    moduli = P.ntt_primes(30, L)
    a, b = rand_lbn(11, moduli, B, N), rand_lbn(12, moduli, B, N)
    R = cref.Ring(N, moduli, [O.find_psi(q, N) for q in moduli])
    ta, tb = dev(a, cuda_device, packed), dev(b, cuda_device, packed)
    want_c, want_X = R.poly_mul(a, b, fast=True, nthreads=16), R.ntt(a, fast=True, nthreads=16)
    with P.ring(moduli):
        for _ in range(2):
            X = P.NTT(ta)
            c = P.poly_mul(ta, tb)
            assert np.array_equal(host(X), want_X)
            assert np.array_equal(host(c), want_c)
            assert torch.equal(P.iNTT(X), ta)
    P.clear_plans()
