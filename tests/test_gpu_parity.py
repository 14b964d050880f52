"""GPU parity tests (run with -m gpu on the B200 box).  Everything goes through the drop-in modules
(arithmetic / primitive / polynomial) or straight through the C ABI; the oracle is the checker only.
Bar: bit-exact.  /root/reference is not read here: its outputs live in tests/golden/*.npz.
"""
import ctypes
import random

import numpy as np
import pytest
import torch

from oracle import cref
from oracle import oracle as O

pytestmark = pytest.mark.gpu


def dev_i64(a: np.ndarray, device) -> torch.Tensor:
    return torch.from_numpy(np.ascontiguousarray(a).view(np.int64)).to(device)


def host_u64(t: torch.Tensor) -> np.ndarray:
    return t.cpu().numpy().view(np.uint64)


def rand_lbn(seed, moduli, B, N):
    rng = np.random.default_rng(seed)
    return np.stack([rng.integers(0, q, (B, N), dtype=np.uint64) for q in moduli])


def oracle_ring(N, moduli):
    return cref.Ring(N, moduli, [O.find_psi(q, N) for q in moduli])


def col(moduli):
    return np.array(list(moduli), dtype=object).reshape(len(moduli), 1, 1)


# ---------------------------------------------------------------------------------------------------
# element-wise ops vs the REFERENCE's own outputs (golden) -- arithmetic.py:3-13
# ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("op", ["vec_add", "vec_sub", "vec_mul"])
def test_vec_ops_match_reference_golden(cuda_device, golden_dir, op):
    import arithmetic
    g = np.load(golden_dir + "/vec_ops.npz")
    for name in g["names"]:
        a, b, mod = g[name + "_a"], g[name + "_b"], g[name + "_mod"]
        want = g[name + "_" + op]
        MOD = col([int(m) for m in mod])
        got = getattr(arithmetic, op)(dev_i64(a, cuda_device), dev_i64(b, cuda_device), MOD)
        assert got.is_cuda and got.dtype == torch.int64 and tuple(got.shape) == a.shape
        assert np.array_equal(host_u64(got), want), (op, name)
        # numpy in -> numpy out (host buffers through the same kernels), object dtype preserved
        got_np = getattr(arithmetic, op)(a.astype(object), b.astype(object), MOD)
        assert got_np.dtype == object and [int(v) for v in got_np.ravel()] == want.ravel().tolist()


def test_vec_ops_shapes_and_edges(cuda_device):
    import arithmetic as A
    q = 1152921504606584833
    # scalar MOD, 1-D, odd length, unaligned views, empty
    for n in (0, 1, 2, 3, 255, 4097):
        rng = np.random.default_rng(n)
        a, b = rng.integers(0, q, n, dtype=np.uint64), rng.integers(0, q, n, dtype=np.uint64)
        for op in ("add", "sub", "mul"):
            got = getattr(A, "vec_" + op)(dev_i64(a, cuda_device), dev_i64(b, cuda_device), q)
            assert np.array_equal(host_u64(got), cref.vec_op(op, a, b, q)), (n, op)
    # a non-contiguous / offset view must not be assumed 16-byte aligned
    base = torch.arange(0, 2 * 1001, device=cuda_device, dtype=torch.int64)
    v = base[1::2]
    got = A.vec_add(v, v, 97)
    assert torch.equal(got, (2 * v) % 97)
    # inputs are never modified (out-of-place like the reference)
    x = dev_i64(np.arange(64, dtype=np.uint64), cuda_device)
    x0 = x.clone()
    A.vec_mul(x, x, 97)
    assert torch.equal(x, x0)
    # uint64 tensors are accepted as containers
    u = x.view(torch.uint64)
    assert A.vec_add(u, u, 97).dtype == torch.uint64
    with pytest.raises(AssertionError):
        A.vec_add(x, x[:-1], 97)


def test_poly_add_returns_both_sums(cuda_device):
    import polynomial as P
    q = 1125899902124033
    rng = np.random.default_rng(5)
    a = [rng.integers(0, q, (2, 64), dtype=np.uint64) for _ in range(2)]
    b = [rng.integers(0, q, (2, 64), dtype=np.uint64) for _ in range(2)]
    ta = tuple(dev_i64(x, cuda_device) for x in a)
    tb = tuple(dev_i64(x, cuda_device) for x in b)
    s = P.poly_add(ta, tb, q)     # " polynomial.py":3-5 drops these; we return them
    d = P.poly_sub(ta, tb, q)
    for i in range(2):
        assert np.array_equal(host_u64(s[i]), cref.vec_op("add", a[i], b[i], q))
        assert np.array_equal(host_u64(d[i]), cref.vec_op("sub", a[i], b[i], q))


# ---------------------------------------------------------------------------------------------------
# transforms vs the definition-level golden vectors and the oracle
# ---------------------------------------------------------------------------------------------------
def test_ntt_matches_definition_golden(cuda_device, golden_dir):
    import arithmetic as A
    g = np.load(golden_dir + "/ntt_small.npz")
    for name in g["names"]:
        q, psi = (int(v) for v in g[name + "_q"])
        x, X = g[name + "_x"], g[name + "_X"]
        got = A.NTT(dev_i64(x, cuda_device), MOD=q)
        assert np.array_equal(host_u64(got), X), name
        back = A.iNTT(got, MOD=q)
        assert np.array_equal(host_u64(back), x), name


@pytest.mark.parametrize("logn", list(range(4, 18)))
def test_ntt_intt_all_sizes(cuda_device, logn):
    """Every kernel instantiation: N = 2^4 .. 2^17, 2 limbs x 3 batch, 60-bit primes."""
    import primitive as P
    N = 1 << logn
    moduli = P.ntt_primes(60, 2)
    x = rand_lbn(logn, moduli, 3, N)
    # extreme residues
    x[:, 0, 0] = 0
    x[0, 0, 1] = moduli[0] - 1
    x[1, 1, :] = moduli[1] - 1
    R = oracle_ring(N, moduli)
    tx = dev_i64(x, cuda_device)
    with P.ring(moduli):
        X = P.NTT(tx)
        back = P.iNTT(X)
    assert np.array_equal(host_u64(X), R.ntt(x, fast=True, nthreads=4))
    assert torch.equal(back, tx)
    assert torch.equal(tx, dev_i64(x, cuda_device))          # input untouched
    # evaluation order, independent of any table: NTT(x)[k] = x(psi^(2 brv(k)+1))
    rng = random.Random(logn)
    Xh = host_u64(X)
    for _ in range(2 if logn > 12 else 4):
        l, b, k = rng.randrange(2), rng.randrange(3), rng.randrange(N)
        assert int(Xh[l, b, k]) == O.ntt_point(x[l, b].tolist(), moduli[l], O.find_psi(moduli[l], N), k)


@pytest.mark.parametrize("logn", [18, 20])
def test_ntt_roundtrip_large(cuda_device, logn):
    import primitive as P
    N = 1 << logn
    moduli = P.ntt_primes(60, 1, N=N)
    x = rand_lbn(logn, moduli, 1, N)
    tx = dev_i64(x, cuda_device)
    with P.ring(moduli):
        X = P.NTT(tx)
        assert torch.equal(P.iNTT(X), tx)
    R = oracle_ring(N, moduli)
    assert np.array_equal(host_u64(X), R.ntt(x, fast=True))


@pytest.mark.parametrize("bits", [30, 50, 61])
def test_modulus_widths(cuda_device, bits):
    """30-bit (config 1), 50-bit (config 2) and the widest supported (61-bit) primes."""
    import primitive as P
    N = 2048
    moduli = P.ntt_primes(bits, 2)
    a, b = rand_lbn(1, moduli, 2, N), rand_lbn(2, moduli, 2, N)
    a[:, :, :8] = np.array(moduli, dtype=np.uint64).reshape(2, 1, 1) - 1
    R = oracle_ring(N, moduli)
    import polynomial as Poly
    c = Poly.poly_mul(dev_i64(a, cuda_device), dev_i64(b, cuda_device), list(moduli))
    assert np.array_equal(host_u64(c), R.poly_mul(a, b, fast=True))


def test_in_place_and_limb_ranges_through_c_abi(cuda_device):
    """Straight C-ABI calls: in == out, and a limb sub-range of a plan (what a limb shard uses)."""
    import primitive as P
    from gpu_fhe_b200 import _capi
    N, B = 8192, 2
    moduli = P.ntt_primes(60, 4)
    plan = P.make_plan(N, moduli, device=cuda_device.index)
    x = rand_lbn(9, moduli, B, N)
    R = oracle_ring(N, moduli)
    want = R.ntt(x)
    t = dev_i64(x, cuda_device)
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = _capi.lib.gfhe_ntt_forward(plan.handle, 0, 4, B, t.data_ptr(), t.data_ptr(), st)       # in place
    assert rc == 0
    assert np.array_equal(host_u64(t), want)
    # limbs [1, 3) only, buffers hold just those two limbs
    sub = dev_i64(x[1:3], cuda_device)
    out = torch.empty_like(sub)
    assert _capi.lib.gfhe_ntt_forward(plan.handle, 1, 2, B, sub.data_ptr(), out.data_ptr(), st) == 0
    assert np.array_equal(host_u64(out), want[1:3])
    back = torch.empty_like(sub)
    assert _capi.lib.gfhe_ntt_inverse(plan.handle, 1, 2, B, out.data_ptr(), back.data_ptr(), st) == 0
    assert torch.equal(back, sub)
    # argument errors
    assert _capi.lib.gfhe_ntt_forward(plan.handle, 3, 2, B, sub.data_ptr(), out.data_ptr(), st) == -7
    need = _capi.lib.gfhe_poly_mul_workspace_bytes(plan.handle, 2, B)
    assert need == 2 * B * N * 8
    assert _capi.lib.gfhe_poly_mul(plan.handle, 1, 2, B, sub.data_ptr(), sub.data_ptr(), out.data_ptr(), None, 0, st) == -8
    assert _capi.lib.gfhe_poly_mul(plan.handle, 1, 2, B, sub.data_ptr(), sub.data_ptr(), sub.data_ptr(), None, 0, st) == -1
    torch.cuda.synchronize()


# ---------------------------------------------------------------------------------------------------
# poly-mul: golden products (schoolbook / Kronecker), properties, BASELINE configs
# ---------------------------------------------------------------------------------------------------
def test_poly_mul_matches_golden_products(cuda_device, golden_dir):
    import polynomial as P
    g = np.load(golden_dir + "/polymul.npz")
    for name in g["names"]:
        q = int(g[name + "_q"][0])
        a, b, c = g[name + "_a"], g[name + "_b"], g[name + "_c"]
        got = P.poly_mul(dev_i64(a, cuda_device), dev_i64(b, cuda_device), q)
        assert np.array_equal(host_u64(got), c), name
        got_np = P.poly_mul(a, b, q)                 # host arrays in, host array out
        assert isinstance(got_np, np.ndarray) and np.array_equal(got_np.view(np.uint64), c), name


@pytest.mark.parametrize("logn", [4, 6, 9, 12, 13, 15, 16, 17])
def test_poly_mul_properties(cuda_device, logn):
    import polynomial as P
    N = 1 << logn
    moduli = P.ntt_primes(60, 2)
    a, b, c2 = rand_lbn(1, moduli, 2, N), rand_lbn(2, moduli, 2, N), rand_lbn(3, moduli, 2, N)
    ta, tb, tc = (dev_i64(v, cuda_device) for v in (a, b, c2))
    MOD = col(moduli)
    with P.ring(moduli):
        ab = P.poly_mul(ta, tb)
        # oracle
        assert np.array_equal(host_u64(ab), oracle_ring(N, moduli).poly_mul(a, b, fast=True, nthreads=4))
        # commutative
        assert torch.equal(ab, P.poly_mul(tb, ta))
        # identity: a * 1 = a
        one = torch.zeros_like(ta)
        one[:, :, 0] = 1
        assert torch.equal(P.poly_mul(ta, one), ta)
        # a * X = negacyclic shift with sign flip
        xp = torch.zeros_like(ta)
        xp[:, :, 1] = 1
        sh = host_u64(P.poly_mul(ta, xp))
        for l, q in enumerate(moduli):
            want = np.concatenate([(q - a[l, :, -1:]) % q, a[l, :, :-1]], axis=1)
            assert np.array_equal(sh[l], want)
        # distributive: a*(b+c) = a*b + a*c
        lhs = P.poly_mul(ta, P.vec_add(tb, tc, MOD))
        rhs = P.vec_add(ab, P.poly_mul(ta, tc), MOD)
        assert torch.equal(lhs, rhs)
        # NTT is linear and multiplicative: NTT(a*b) = NTT(a) . NTT(b)
        assert torch.equal(P.NTT(ab), P.vec_mul(P.NTT(ta), P.NTT(tb), MOD))


def test_config1_n1024_30bit(cuda_device):
    """BASELINE config 1: N=1024 NTT->INTT round trip + poly-mul, one 30-bit prime, [1,1,1024]."""
    import polynomial as P
    q, N = 1073479681, 1024
    assert P.find_psi(q, N) == 530309095
    rng = np.random.default_rng(20261104 + 1)
    a, b = rng.integers(0, q, N, dtype=np.uint64), rng.integers(0, q, N, dtype=np.uint64)
    ta, tb = dev_i64(a, cuda_device), dev_i64(b, cuda_device)
    X = P.NTT(ta, MOD=q)
    assert host_u64(X).tolist() == O.ntt_definition(a.tolist(), q, 530309095)
    assert torch.equal(P.iNTT(X, MOD=q), ta)
    c = P.poly_mul(ta, tb, q)
    assert host_u64(c).tolist() == O.schoolbook_negacyclic(a.tolist(), b.tolist(), q)


def test_config2_n4096_4x50bit_batch256(cuda_device):
    """BASELINE config 2: [4,256,4096], 50-bit primes -- full-array comparison."""
    import polynomial as P
    N, B = 4096, 256
    moduli = P.ntt_primes(50, 4)
    assert moduli[0] == 1125899902124033
    a, b = rand_lbn(20261104 + 2, moduli, B, N), rand_lbn(99, moduli, B, N)
    c = P.poly_mul(dev_i64(a, cuda_device), dev_i64(b, cuda_device), list(moduli))
    assert np.array_equal(host_u64(c), oracle_ring(N, moduli).poly_mul(a, b, fast=True, nthreads=16))
    # independent of every NTT: Kronecker substitution on a sample
    ch = host_u64(c)
    for (l, bi) in [(0, 0), (3, 255)]:
        assert ch[l, bi].tolist() == O.kronecker_negacyclic(a[l, bi].tolist(), b[l, bi].tolist(), moduli[l])


def test_config3_n16384_8limbs_batch128(cuda_device):
    """BASELINE config 3: [8,128,16384] NTT and INTT."""
    import primitive as P
    N, B = 16384, 128
    moduli = P.ntt_primes(60, 8)
    x = rand_lbn(20261104 + 3, moduli, B, N)
    tx = dev_i64(x, cuda_device)
    with P.ring(moduli):
        X = P.NTT(tx)
        back = P.iNTT(X)
    assert torch.equal(back, tx)
    assert np.array_equal(host_u64(X), oracle_ring(N, moduli).ntt(x, fast=True, nthreads=16))


def test_config4_n65536_32x60bit_batch16(cuda_device):
    """BASELINE config 4 at full size: [32,16,65536] poly-mul, 60-bit primes.

    Full-array comparison against the C oracle (fast flavour, itself pinned to the exact flavour and
    to Kronecker in the CPU suite), plus size-independent checks: one Kronecker product, the paired
    point-evaluation / schoolbook-coefficient spot checks on every limb."""
    import polynomial as P
    N, L, B = 65536, 32, 16
    moduli = P.ntt_primes(60, L)
    assert moduli[0] == 1152921504606584833
    a, b = rand_lbn(20261104 + 4, moduli, B, N), rand_lbn(77, moduli, B, N)
    ta, tb = dev_i64(a, cuda_device), dev_i64(b, cuda_device)
    c = P.poly_mul(ta, tb, list(moduli))
    ch = host_u64(c)
    want = oracle_ring(N, moduli).poly_mul(a, b, fast=True, nthreads=32)
    assert np.array_equal(ch, want)
    assert ch[0, 0].tolist() == O.kronecker_negacyclic(a[0, 0].tolist(), b[0, 0].tolist(), moduli[0])
    rng = random.Random(4)
    for l in range(L):
        bi, k = rng.randrange(B), rng.randrange(N)
        psi = O.find_psi(moduli[l], N)
        al, bl, cl = a[l, bi].tolist(), b[l, bi].tolist(), ch[l, bi].tolist()
        assert O.check_point_eval(al, bl, cl, moduli[l], psi, k)
        assert O.schoolbook_coeff(al, bl, moduli[l], k) == cl[k]


def test_host_entry_points(cuda_device):
    """gfhe_poly_mul_host / gfhe_ntt_host: host arrays, copies pipelined over limb groups."""
    import arithmetic  # noqa: F401
    import primitive as P
    from gpu_fhe_b200 import dispatch
    N, L, B = 16384, 6, 40            # 5 MiB per limb -> several limb groups, last one ragged
    moduli = P.ntt_primes(60, L)
    a, b = rand_lbn(1, moduli, B, N), rand_lbn(2, moduli, B, N)
    R = oracle_ring(N, moduli)
    out = np.empty_like(a)
    dispatch.poly_mul_host(a, b, out, moduli, device=cuda_device.index)
    assert np.array_equal(out, R.poly_mul(a, b, fast=True, nthreads=16))
    X = np.empty_like(a)
    dispatch.ntt_host(a, X, moduli, device=cuda_device.index)
    assert np.array_equal(X, R.ntt(a, fast=True, nthreads=16))
    back = np.empty_like(a)
    dispatch.ntt_host(X, back, moduli, inverse=True, device=cuda_device.index)
    assert np.array_equal(back, a)
    # pinned host memory path
    pa = torch.from_numpy(a.view(np.int64)).pin_memory()
    pb = torch.from_numpy(b.view(np.int64)).pin_memory()
    pc = torch.empty_like(pa).pin_memory()
    dispatch.poly_mul_host(pa.numpy().view(np.uint64), pb.numpy().view(np.uint64), pc.numpy().view(np.uint64), moduli,
                           device=cuda_device.index)
    assert np.array_equal(pc.numpy().view(np.uint64), out)


def test_sharded_equals_unsharded(cuda_device):
    """Limb/batch shards computed independently (as each rank would) reassemble to the full result."""
    import polynomial as P
    from gpu_fhe_b200 import shard
    N, L, B = 8192, 4, 4
    moduli = P.ntt_primes(60, L)
    a, b = rand_lbn(1, moduli, B, N), rand_lbn(2, moduli, B, N)
    ta, tb = dev_i64(a, cuda_device), dev_i64(b, cuda_device)
    full = P.poly_mul(ta, tb, list(moduli))
    for world in (2, 4, 8):
        out = torch.empty_like(full)
        for r in range(world):
            sh = shard.partition(L, B, world, r)
            part = P.poly_mul(shard.take(ta, sh), shard.take(tb, sh), list(moduli[sh.limbs[0]:sh.limbs[1]]))
            out[sh.limbs[0]:sh.limbs[1], sh.batch[0]:sh.batch[1]] = part
        assert torch.equal(out, full), world


# ---------------------------------------------------------------------------------------------------
# NTT-domain-resident ops: inner product and Galois automorphism
# ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("bits,K", [(30, 1), (50, 3), (60, 7), (60, 256), (61, 16)])
def test_ntt_mul_acc_matches_python_ints(cuda_device, bits, K):
    import polynomial as P
    L, B, N = 3, 2, 64
    moduli = O.ntt_primes(bits, L)
    rng = np.random.default_rng(bits * 1000 + K)
    a = np.stack([np.stack([rng.integers(0, q, (B, N), dtype=np.uint64) for q in moduli]) for _ in range(K)])
    b = np.stack([np.stack([rng.integers(0, q, (B, N), dtype=np.uint64) for q in moduli]) for _ in range(K)])
    a[0, :, 0, :4] = np.array([[q - 1] * 4 for q in moduli], dtype=np.uint64)   # extremes: (q-1)^2 in every term slot 0
    b[0, :, 0, :4] = np.array([[q - 1] * 4 for q in moduli], dtype=np.uint64)
    got = host_u64(P.poly_mac_ntt(dev_i64(a, cuda_device), dev_i64(b, cuda_device), MOD=moduli))
    assert got.shape == (L, B, N)
    for l, q in enumerate(moduli):
        want = O.mul_acc([a[t, l].ravel().tolist() for t in range(K)], [b[t, l].ravel().tolist() for t in range(K)], q)
        assert got[l].ravel().tolist() == want, (bits, K, l)
    if K == 1:   # un-stacked form == vec_mul on canonical residues
        import arithmetic as A
        g1 = P.poly_mul_ntt(dev_i64(a[0], cuda_device), dev_i64(b[0], cuda_device), MOD=moduli)
        g2 = A.vec_mul(dev_i64(a[0], cuda_device), dev_i64(b[0], cuda_device), col(moduli))
        assert torch.equal(g1, g2) and np.array_equal(host_u64(g1), got)


def test_ntt_mul_acc_too_many_terms_is_an_error(cuda_device):
    import polynomial as P
    from gpu_fhe_b200 import _capi
    q = O.ntt_primes(61, 1)[0]   # 61-bit modulus: 2^(128-122) = 64 terms at most
    x = torch.zeros((65, 1, 1, 16), dtype=torch.int64, device=cuda_device)
    with pytest.raises(_capi.GpuFheError) as ei:
        P.poly_mac_ntt(x, x, MOD=[q])
    assert ei.value.code == -10


def test_poly_mul_equals_intt_of_ntt_domain_product(cuda_device):
    import polynomial as P
    N, L, B = 8192, 2, 3
    moduli = O.ntt_primes(60, L)
    a, b = rand_lbn(11, moduli, B, N), rand_lbn(12, moduli, B, N)
    ta, tb = dev_i64(a, cuda_device), dev_i64(b, cuda_device)
    with P.ring(moduli):
        c = P.poly_mul(ta, tb)
        c2 = P.iNTT(P.poly_mul_ntt(P.NTT(ta), P.NTT(tb)))
        # inner product of two terms == sum of the two products
        s = P.iNTT(P.poly_mac_ntt([P.NTT(ta), P.NTT(tb)], [P.NTT(tb), P.NTT(tb)]))
        want = P.vec_add(P.poly_mul(ta, tb), P.poly_mul(tb, tb), col(moduli))
    assert torch.equal(c, c2)
    assert torch.equal(s, want)


@pytest.mark.parametrize("logn", [4, 8, 12, 16])
def test_automorphism_both_domains(cuda_device, logn):
    import polynomial as P
    N, L, B = 1 << logn, 2, 2
    moduli = O.ntt_primes(60, L)
    a = rand_lbn(logn, moduli, B, N)
    a[:, :, 1] = 0   # zero coefficients must stay zero under negation (not become q)
    ta = dev_i64(a, cuda_device)
    for k in (3, 5, 2 * N - 1, N + 1, 1, 2 * N + 3):
        with P.ring(moduli):
            c = P.automorphism(ta, k)
            cn = P.automorphism(P.NTT(ta), k, ntt=True)
            assert torch.equal(P.NTT(c), cn), (logn, k)          # commutes with the transform
        h = host_u64(c)
        for l, q in enumerate(moduli):
            want = O.automorphism_coeff(a[l, 0].tolist(), k % (2 * N), q)
            assert h[l, 0].tolist() == want, (logn, k, l)
    if logn <= 8:   # NTT-domain permutation against the definition-level oracle
        q = moduli[0]
        psi = O.find_psi(q, N)
        X = O.ntt_definition(a[0, 0].tolist(), q, psi)
        with P.ring(moduli):
            got = host_u64(P.automorphism(P.NTT(ta), 5, ntt=True))[0, 0].tolist()
        assert got == O.automorphism_ntt(X, 5)
    with pytest.raises(ValueError):
        P.automorphism(ta, 4, MOD=moduli)


def test_poly_mul_pair_is_the_tensor_product(cuda_device):
    import polynomial as P
    N, L, B = 4096, 2, 2
    moduli = O.ntt_primes(50, L)
    xs = [dev_i64(rand_lbn(40 + i, moduli, B, N), cuda_device) for i in range(4)]
    M = col(moduli)
    with P.ring(moduli):
        c0, c1, c2 = P.poly_mul_pair((xs[0], xs[1]), (xs[2], xs[3]))
        assert torch.equal(c0, P.poly_mul(xs[0], xs[2]))
        assert torch.equal(c2, P.poly_mul(xs[1], xs[3]))
        assert torch.equal(c1, P.vec_add(P.poly_mul(xs[0], xs[3]), P.poly_mul(xs[1], xs[2]), M))
    # one limb-polynomial of c1 against the schoolbook oracle
    q = moduli[0]
    a0, a1, b0, b1 = (host_u64(x)[0, 0].tolist() for x in xs)
    want = O.vec_add(O.kronecker_negacyclic(a0, b1, q), O.kronecker_negacyclic(a1, b0, q), q)
    assert host_u64(c1)[0, 0].tolist() == want


# ---------------------------------------------------------------------------------------------------
# 32-bit arithmetic policy (moduli below 2^30): every kernel instantiation against the C oracle, and
# bit-identical to the 64-bit kernels on the same inputs
# ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("logn", [4, 5, 8, 10, 12, 13, 14, 16, 17, 18])
def test_w32_policy_all_sizes(cuda_device, logn, monkeypatch):
    import polynomial as P
    N = 1 << logn
    moduli = P.ntt_primes(30, 2, N=max(N, 1 << 12))
    assert all(q < (1 << 30) for q in moduli)
    B = 3 if logn <= 16 else 1
    a, b = rand_lbn(100 + logn, moduli, B, N), rand_lbn(200 + logn, moduli, B, N)
    a[:, 0, :4] = np.array(moduli, dtype=np.uint64).reshape(2, 1) - 1      # extremes: q-1 and 0
    b[:, 0, :4] = np.array(moduli, dtype=np.uint64).reshape(2, 1) - 1
    a[:, 0, 4] = 0
    R = oracle_ring(N, moduli)
    ta, tb = dev_i64(a, cuda_device), dev_i64(b, cuda_device)
    P.clear_plans()
    with P.ring(moduli):
        X, c = P.NTT(ta), P.poly_mul(ta, tb)
        back = P.iNTT(X)
    assert np.array_equal(host_u64(X), R.ntt(a, fast=True, nthreads=4))
    assert torch.equal(back, ta)
    assert np.array_equal(host_u64(c), R.poly_mul(a, b, fast=True, nthreads=4))
    # the same plan forced onto the 64-bit kernels gives identical words
    monkeypatch.setenv("GFHE_NO_W32", "1")
    P.clear_plans()
    with P.ring(moduli):
        assert torch.equal(P.NTT(ta), X) and torch.equal(P.poly_mul(ta, tb), c)
    monkeypatch.delenv("GFHE_NO_W32")
    P.clear_plans()


@pytest.mark.parametrize("logn,bits", [(4, 60), (6, 30), (10, 50), (12, 60), (13, 61), (14, 30), (16, 60), (17, 60)])
def test_poly_mul_pre_equals_poly_mul(cuda_device, logn, bits):
    """One operand cached in the NTT domain: same words as the coefficient-domain product, every kernel shape/policy."""
    import polynomial as P
    N, L, B = 1 << logn, 2, 3
    moduli = P.ntt_primes(bits, L)
    a, b = rand_lbn(300 + logn, moduli, B, N), rand_lbn(400 + logn, moduli, B, N)
    a[:, 0, :3] = np.array(moduli, dtype=np.uint64).reshape(2, 1) - 1
    b[:, 0, :3] = np.array(moduli, dtype=np.uint64).reshape(2, 1) - 1
    ta, tb = dev_i64(a, cuda_device), dev_i64(b, cuda_device)
    with P.ring(moduli):
        Bn = P.NTT(tb)
        c = P.poly_mul_pre(ta, Bn)
        assert torch.equal(c, P.poly_mul(ta, tb))
        assert torch.equal(Bn, P.NTT(tb)) and torch.equal(ta, dev_i64(a, cuda_device))   # inputs untouched
    if logn <= 13:
        R = oracle_ring(N, moduli)
        assert np.array_equal(host_u64(c), R.poly_mul(a, b, fast=True))


# ---------------------------------------------------------------------------------------------------
# FP64 arithmetic policy (moduli below 2^50): exact by construction (error-free products), so it must give the same
# words as the integer kernels and the C oracle for every size, modulus width and extreme input
# ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("logn,bits", [(4, 50), (5, 50), (8, 31), (10, 50), (12, 50), (12, 40), (13, 50), (14, 45),
                                       (16, 50), (17, 50), (18, 50)])
def test_fp64_policy_all_sizes(cuda_device, logn, bits, monkeypatch):
    import polynomial as P
    N = 1 << logn
    moduli = P.ntt_primes(bits, 2, N=max(N, 1 << 12))
    assert all((1 << 30) <= q < (1 << 50) for q in moduli)
    B = 3 if logn <= 16 else 1
    a, b = rand_lbn(500 + logn, moduli, B, N), rand_lbn(600 + logn, moduli, B, N)
    qm1 = np.array(moduli, dtype=np.uint64).reshape(2, 1) - 1
    a[:, 0, :4] = qm1
    b[:, 0, :4] = qm1
    a[:, 0, 4] = 0
    if B > 1:
        a[:, 1, :] = qm1            # a whole polynomial of q-1 against random and against itself
        b[:, 2, :] = a[:, 2, :]
    R = oracle_ring(N, moduli)
    ta, tb = dev_i64(a, cuda_device), dev_i64(b, cuda_device)
    monkeypatch.delenv("GFHE_NO_F64", raising=False)
    P.clear_plans()
    with P.ring(moduli):
        X, c = P.NTT(ta), P.poly_mul(ta, tb)
        back = P.iNTT(X)
        pre = P.poly_mul_pre(ta, P.NTT(tb))
    assert np.array_equal(host_u64(X), R.ntt(a, fast=True, nthreads=4))
    assert torch.equal(back, ta)
    assert np.array_equal(host_u64(c), R.poly_mul(a, b, fast=True, nthreads=4))
    assert torch.equal(pre, c)
    monkeypatch.setenv("GFHE_NO_F64", "1")
    P.clear_plans()
    with P.ring(moduli):
        assert torch.equal(P.NTT(ta), X) and torch.equal(P.poly_mul(ta, tb), c) and torch.equal(P.iNTT(X), ta)
    monkeypatch.delenv("GFHE_NO_F64")
    P.clear_plans()


@pytest.mark.parametrize("bits", [30, 50, 60, 61])
@pytest.mark.parametrize("logn", [13, 16, 17])
def test_persistent_row_pass_matches_one_tile_per_cta(cuda_device, logn, bits, monkeypatch):
    """Two-pass products: the persistent fused row pass (TMA prefetch of the next tile, mbarrier phases per trip) and the
    one-tile-per-CTA kernel give the same words, for every policy, with more tiles than resident CTAs (several trips
    per CTA, both mbarrier parities) and with fewer; checked against the C oracle at 2^13."""
    import polynomial as P
    N, L = 1 << logn, 2
    B = 40 if logn == 16 else 3          # 2 * 40 * 16 = 1280 tiles: 4-5 trips per resident CTA
    moduli = P.ntt_primes(bits, L, N=max(N, 1 << 12))
    a, b = rand_lbn(700 + logn, moduli, B, N), rand_lbn(800 + logn, moduli, B, N)
    qm1 = np.array(moduli, dtype=np.uint64).reshape(L, 1) - 1
    a[:, 0, :4] = qm1
    b[:, 0, :4] = qm1
    a[:, 1, :] = qm1
    ta, tb = dev_i64(a, cuda_device), dev_i64(b, cuda_device)
    out = {}
    for mode in ("0", "1"):
        monkeypatch.setenv("GFHE_PERSIST", mode)
        P.clear_plans()
        with P.ring(moduli):
            out[mode] = P.poly_mul(ta, tb)
    monkeypatch.delenv("GFHE_PERSIST")
    P.clear_plans()
    assert torch.equal(out["0"], out["1"])
    with P.ring(moduli):
        assert torch.equal(P.poly_mul(ta, tb), out["0"])      # the per-policy default
    if logn == 13:
        R = oracle_ring(N, moduli)
        assert np.array_equal(host_u64(out["1"]), R.poly_mul(a, b, fast=True, nthreads=4))


# ---------------------------------------------------------------------------------------------------
# packed 32-bit layout (int32 / uint32 tensors, every modulus below 2^30): same residues as the 64-bit layout
# ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("logn", [4, 5, 8, 10, 12, 13, 14, 16, 17, 18])
def test_packed_u32_layout_all_sizes(cuda_device, logn):
    import polynomial as P
    N, L = 1 << logn, 2
    moduli = P.ntt_primes(30, L, N=max(N, 1 << 12))
    B = 40 if logn == 16 else (3 if logn < 16 else 1)      # 2^16: several trips of the persistent row pass
    a, b = rand_lbn(900 + logn, moduli, B, N), rand_lbn(950 + logn, moduli, B, N)
    qm1 = np.array(moduli, dtype=np.uint64).reshape(L, 1) - 1
    a[:, 0, :4] = qm1
    b[:, 0, :4] = qm1
    a[:, 0, 4] = 0
    ta, tb = dev_i64(a, cuda_device), dev_i64(b, cuda_device)
    pa, pb = ta.to(torch.int32), tb.to(torch.int32)
    with P.ring(moduli):
        X64, c64 = P.NTT(ta), P.poly_mul(ta, tb)
        X, c = P.NTT(pa), P.poly_mul(pa, pb)
        back = P.iNTT(X)
        pre = P.poly_mul_pre(pa, P.NTT(pb))
    assert X.dtype == c.dtype == back.dtype == torch.int32 and X.shape == pa.shape
    assert torch.equal(X.to(torch.int64), X64)
    assert torch.equal(c.to(torch.int64), c64)
    assert torch.equal(back, pa) and torch.equal(pre, c)
    assert torch.equal(pa, ta.to(torch.int32)) and torch.equal(pb, tb.to(torch.int32))       # inputs untouched
    if logn <= 14:
        R = oracle_ring(N, moduli)
        assert np.array_equal(X.to(torch.int64).cpu().numpy().view(np.uint64), R.ntt(a, fast=True, nthreads=4))
        assert np.array_equal(c.to(torch.int64).cpu().numpy().view(np.uint64), R.poly_mul(a, b, fast=True, nthreads=4))


def test_packed_u32_layout_errors(cuda_device):
    import polynomial as P
    from gpu_fhe_b200 import _capi
    N = 256
    x = torch.zeros((2, 1, N), dtype=torch.int32, device=cuda_device)
    with P.ring(P.ntt_primes(50, 2)):
        with pytest.raises(_capi.GpuFheError) as e:
            P.NTT(x)
        assert e.value.code == -12                                   # GFHE_ELAYOUT: a modulus does not fit 30 bits
    m30 = P.ntt_primes(30, 2)
    with P.ring(m30):
        with pytest.raises(TypeError):
            P.poly_mul(x, x.to(torch.int64))                         # mixed word sizes
    with pytest.raises(_capi.GpuFheError) as e:
        P.vec_add(x, x, np.array(P.ntt_primes(50, 2), dtype=np.int64).reshape(2, 1, 1))
    assert e.value.code == -12


def test_packed_u32_elementwise_mac_automorphism(cuda_device):
    """The operations around the transforms on int32 tensors: same residues as the int64 kernels (which are pinned to
    the reference's outputs and the oracle), for ANY int32 inputs of the element-wise ops."""
    import polynomial as P
    N, L, B = 1 << 10, 3, 5
    moduli = P.ntt_primes(30, L, N=1 << 12)
    MOD = np.array(moduli, dtype=np.int64).reshape(L, 1, 1)
    rng = np.random.default_rng(77)
    a = rng.integers(-2**31, 2**31, (L, B, N), dtype=np.int64)
    b = rng.integers(-2**31, 2**31, (L, B, N), dtype=np.int64)
    a[:, 0, :4] = [-2**31, 2**31 - 1, 0, -1]
    b[:, 0, :4] = [-2**31, 2**31 - 1, -1, 2**31 - 1]
    ta, tb = torch.from_numpy(a).to(cuda_device), torch.from_numpy(b).to(cuda_device)
    pa, pb = ta.to(torch.int32), tb.to(torch.int32)
    ao, bo, mo = a.astype(object), b.astype(object), MOD.astype(object)
    for op, want in ((P.vec_add, (ao + bo) % mo), (P.vec_sub, (ao - bo) % mo), (P.vec_mul, (ao * bo) % mo)):   # arithmetic.py:5,9,13 on Python ints
        got = op(pa, pb, MOD)
        assert got.dtype == torch.int32
        assert torch.equal(got.to(torch.int64), op(ta, tb, MOD))
        assert np.array_equal(got.cpu().numpy().astype(object), want)
    odd = op(pa[:, :, :1001].contiguous(), pb[:, :, :1001].contiguous(), MOD)
    assert torch.equal(odd.to(torch.int64), P.vec_mul(ta[:, :, :1001].contiguous(), tb[:, :, :1001].contiguous(), MOD))
    # canonical residues for the NTT-domain ops
    ra, rb = rand_lbn(71, moduli, B, N), rand_lbn(72, moduli, B, N)
    ua, ub = dev_i64(ra, cuda_device), dev_i64(rb, cuda_device)
    with P.ring(moduli):
        for K in (1, 3, 8, 19):
            sa = torch.stack([(ua + 7 * t) % torch.from_numpy(MOD).to(cuda_device) for t in range(K)])
            sb = torch.stack([(ub + 11 * t) % torch.from_numpy(MOD).to(cuda_device) for t in range(K)])
            want = P.poly_mac_ntt(list(sa), list(sb))
            got = P.poly_mac_ntt(list(sa.to(torch.int32)), list(sb.to(torch.int32)))
            assert got.dtype == torch.int32 and torch.equal(got.to(torch.int64), want)
        for k in (3, 5, 2 * N - 1, 1025):
            for ntt in (False, True):
                got = P.automorphism(ua.to(torch.int32), k, ntt=ntt)
                assert torch.equal(got.to(torch.int64), P.automorphism(ua, k, ntt=ntt))




def test_empty_and_non_contiguous_operands(cuda_device):
    """Edge cases of the marshalling layer: zero-sized batches come back empty, strided views are made contiguous,
    inputs are never written."""
    import polynomial as P
    N, L = 256, 2
    moduli = P.ntt_primes(60, L)
    empty = torch.empty((L, 0, N), dtype=torch.int64, device=cuda_device)
    with P.ring(moduli):
        assert tuple(P.NTT(empty).shape) == (L, 0, N)
        assert tuple(P.iNTT(empty).shape) == (L, 0, N)
        assert tuple(P.poly_mul(empty, empty).shape) == (L, 0, N)
        assert tuple(P.poly_mul_pre(empty, empty).shape) == (L, 0, N)
        assert tuple(P.poly_mul_ntt(empty, empty).shape) == (L, 0, N)
        assert tuple(P.automorphism(empty, 3).shape) == (L, 0, N)
    big = dev_i64(rand_lbn(77, moduli, 4, 2 * N), cuda_device)
    view = big[:, ::2, :N]                     # non-contiguous in two dimensions
    assert not view.is_contiguous()
    keep = big.clone()
    with P.ring(moduli):
        X = P.NTT(view)
        assert torch.equal(X, P.NTT(view.contiguous()))
        assert torch.equal(P.iNTT(X), view)
        c = P.poly_mul(view, view)
        assert torch.equal(c, P.poly_mul(view.contiguous(), view.contiguous()))
    assert torch.equal(big, keep)
    # [B, N] and [N] operands with a single modulus
    q = moduli[0]
    x = dev_i64(rand_lbn(78, [q], 1, N)[0, 0], cuda_device)
    assert tuple(P.NTT(x, MOD=q).shape) == (N,)
    assert torch.equal(P.iNTT(P.NTT(x, MOD=q), MOD=q), x)
    with pytest.raises(ValueError):
        P.NTT(big, MOD=[q, q, q])              # three moduli, leading dimension 2
    with pytest.raises(RuntimeError):
        P.set_ring(None); P.NTT(x)             # no ring installed and no MOD=


@pytest.mark.parametrize("L,B", [(1, 1), (64, 64)])
def test_config5_extremes_n131072(cuda_device, L, B):
    """BASELINE.json configs[4] end points: N = 2^17 forward NTT at [1,1] and [64,64] (4096 limb-NTTs, 4 GiB per buffer), every
    word against the C oracle; the inverse must give the input back."""
    import polynomial as P
    N = 1 << 17
    moduli = P.ntt_primes(60, L)
    rng = np.random.default_rng(20261104 + 5)
    x = np.empty((L, B, N), dtype=np.uint64)
    for l, q in enumerate(moduli):
        x[l] = rng.integers(0, q, (B, N), dtype=np.uint64)
    x[:, 0, :4] = np.array(moduli, dtype=np.uint64).reshape(L, 1) - 1
    tx = dev_i64(x, cuda_device)
    with P.ring(moduli):
        X = P.NTT(tx)
        back = P.iNTT(X)
    assert torch.equal(back, tx)
    del back
    got = host_u64(X)
    del X
    want = oracle_ring(N, moduli).ntt(x, fast=True, nthreads=32)
    assert np.array_equal(got, want)
