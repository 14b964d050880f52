"""GPU tests of the drop-in contract around the kernels (run with -m gpu): numpy broadcasting of MOD, containers and
their value ranges, buffer alignment, the serialisable plan cache, device handling, the pageable host path, and the
N-GPU sharded product with its all-gather.  The oracle (and, for vec_*, the reference's own statement `(a op b) % MOD`
on dtype=object arrays, /root/reference/arithmetic.py:5,9,13) is the checker only.
"""
import ctypes
import os
import socket

import numpy as np
import pytest
import torch

from oracle import cref
from oracle import oracle as O

pytestmark = pytest.mark.gpu

OPS = {"vec_add": lambda a, b: a + b, "vec_sub": lambda a, b: a - b, "vec_mul": lambda a, b: a * b}


def dev_i64(a: np.ndarray, device) -> torch.Tensor:
    return torch.from_numpy(np.ascontiguousarray(a).view(np.int64)).to(device)


def host_u64(t: torch.Tensor) -> np.ndarray:
    return t.cpu().numpy().view(np.uint64)


def rand_lbn(seed, moduli, B, N):
    rng = np.random.default_rng(seed)
    return np.stack([rng.integers(0, q, (B, N), dtype=np.uint64) for q in moduli])


def oracle_ring(N, moduli):
    return cref.Ring(N, moduli, [O.find_psi(q, N) for q in moduli])


# ---------------------------------------------------------------------------------------------------
# vec_*: MOD is anything numpy broadcasts (arithmetic.py:5 `(a + b) % MOD`)
# ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("op", sorted(OPS))
@pytest.mark.parametrize("mod_shape", [(), (3, 1, 1), (8,), (4, 1), (1, 4, 8), (3, 4, 8), (2, 1, 1, 1), (3, 1, 8)])
def test_vec_mod_broadcasting_follows_numpy(cuda_device, op, mod_shape):
    import arithmetic as A
    L, B, N = 3, 4, 8
    rng = np.random.default_rng(1000 * len(op) + 17 * len(mod_shape) + sum(mod_shape))
    primes = O.ntt_primes(60, 8) + O.ntt_primes(50, 8) + O.ntt_primes(30, 8) + [97, 2, 3, (1 << 62) - 57]
    cnt = int(np.prod(mod_shape)) if mod_shape else 1
    mod = np.array([primes[i % len(primes)] for i in rng.integers(0, len(primes), cnt)], dtype=object).reshape(mod_shape)
    a = rng.integers(-(1 << 63), (1 << 63) - 1, (L, B, N), dtype=np.int64)    # any int64, negatives included
    b = rng.integers(-(1 << 63), (1 << 63) - 1, (L, B, N), dtype=np.int64)
    want = OPS[op](a.astype(object), b.astype(object)) % mod                    # the reference's statement, exact mode
    got = getattr(A, op)(dev_i64(a, cuda_device), dev_i64(b, cuda_device), mod if mod_shape else int(mod))
    assert tuple(got.shape) == want.shape
    assert np.array_equal(got.cpu().numpy().astype(object), want), (op, mod_shape)


def test_vec_mod_does_not_broadcast_raises_like_numpy(cuda_device):
    import arithmetic as A
    x = torch.zeros((2, 3, 4), dtype=torch.int64, device=cuda_device)
    with pytest.raises(ValueError):
        A.vec_add(x, x, np.array([5, 7, 11]))


def test_containers_and_value_ranges(cuda_device):
    import arithmetic as A
    q = 1152921504606584833
    big = np.array([(1 << 64) - 1, (1 << 63), (1 << 63) + 12345, 7], dtype=np.uint64)
    # uint64 containers: values >= 2^63 are large positive numbers, not negative ones
    tu = torch.from_numpy(big.view(np.int64)).to(cuda_device).view(torch.uint64)
    for op, f in OPS.items():
        want = f(big.astype(object), big.astype(object)[::-1]) % q
        got = getattr(A, op)(tu, tu.view(torch.int64).flip(0).view(torch.uint64), q)
        assert got.dtype == torch.uint64
        assert np.array_equal(got.cpu().numpy().astype(object), want), op
        got_np = getattr(A, op)(big, big[::-1].copy(), q)                      # numpy uint64 in -> numpy uint64 out
        assert got_np.dtype == np.uint64 and np.array_equal(got_np.astype(object), want)
    # the same words in an int64 container ARE negative numbers (Python % semantics)
    ti = tu.view(torch.int64)
    want = (big.view(np.int64).astype(object) * 3) % q
    assert np.array_equal(A.vec_mul(ti, torch.full_like(ti, 3), q).cpu().numpy().astype(object), want)
    # dtype=object: exact Python ints, signed or unsigned reading chosen by the values
    obj_neg = np.array([-5, (1 << 63) - 1, -(1 << 63)], dtype=object)
    assert np.array_equal(A.vec_add(obj_neg, obj_neg, q), (obj_neg + obj_neg) % q)
    obj_big = np.array([(1 << 64) - 1, 1 << 63, 0], dtype=object)
    r = A.vec_mul(obj_big, obj_big, q)
    assert r.dtype == object and np.array_equal(r, (obj_big * obj_big) % q)
    with pytest.raises(ValueError, match="64-bit word"):
        A.vec_add(np.array([1 << 64], dtype=object), np.array([0], dtype=object), q)
    # mixed containers: fine while the signed one holds no negative value
    assert np.array_equal(A.vec_add(tu, torch.ones_like(ti), q).cpu().numpy().astype(object), (big.astype(object) + 1) % q)
    with pytest.raises(ValueError, match="negative"):
        A.vec_add(tu, -torch.ones_like(ti), q)


def test_many_moduli_rows(cuda_device):
    """One modulus per coefficient (MOD of shape [N] with N > 65535 rows: more than one grid.y span)."""
    import arithmetic as A
    n = 70001
    rng = np.random.default_rng(3)
    mod = rng.integers(2, 1 << 61, n, dtype=np.int64) | 1
    a = rng.integers(-(1 << 62), 1 << 62, (2, n), dtype=np.int64)
    got = A.vec_mul(dev_i64(a, cuda_device), dev_i64(a[::-1].copy(), cuda_device), mod)
    want = (a.astype(object) * a[::-1].astype(object)) % mod.astype(object)
    assert np.array_equal(got.cpu().numpy().astype(object), want)


# ---------------------------------------------------------------------------------------------------
# alignment: the transform kernels use 16-byte accesses
# ---------------------------------------------------------------------------------------------------
def test_misaligned_contiguous_view(cuda_device):
    import polynomial as P
    from gpu_fhe_b200 import _capi
    N, L, B = 1024, 2, 3
    moduli = P.ntt_primes(60, L)
    a, b = rand_lbn(1, moduli, B, N), rand_lbn(2, moduli, B, N)
    R = oracle_ring(N, moduli)
    flat_a = torch.zeros(L * B * N + 2, dtype=torch.int64, device=cuda_device)
    flat_b = torch.zeros_like(flat_a)
    va = flat_a[1:1 + L * B * N].view(L, B, N)          # contiguous, but 8 bytes past a 16-byte boundary
    vb = flat_b[1:1 + L * B * N].view(L, B, N)
    va.copy_(dev_i64(a, cuda_device)); vb.copy_(dev_i64(b, cuda_device))
    assert va.is_contiguous() and va.data_ptr() % 16 == 8
    with P.ring(moduli):
        assert np.array_equal(host_u64(P.NTT(va)), R.ntt(a))
        assert torch.equal(P.iNTT(P.NTT(va)), va)
        assert np.array_equal(host_u64(P.poly_mul(va, vb)), R.poly_mul(a, b))
        assert np.array_equal(host_u64(P.poly_mul_pre(va, P.NTT(vb))), R.poly_mul(a, b))
    # straight through the C ABI such a pointer is an argument error, not a sticky misaligned-address fault
    plan = P.make_plan(N, moduli, device=cuda_device.index)
    out = torch.empty((L, B, N), dtype=torch.int64, device=cuda_device)
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    lib = _capi.lib
    assert lib.gfhe_ntt_forward(plan.handle, 0, L, B, va.data_ptr(), out.data_ptr(), st) == -13
    assert lib.gfhe_ntt_inverse(plan.handle, 0, L, B, out.data_ptr(), va.data_ptr(), st) == -13
    assert lib.gfhe_poly_mul(plan.handle, 0, L, B, va.data_ptr(), out.data_ptr(), out.data_ptr() + 16, None, 0, st) == -13
    assert lib.gfhe_poly_mul_pre(plan.handle, 0, L, B, out.data_ptr(), vb.data_ptr(), va.data_ptr(), st) == -13
    torch.cuda.synchronize()
    assert torch.equal(P.iNTT(P.NTT(va, MOD=moduli), MOD=moduli), va)     # the context is still healthy
    # packed layout: a 4-byte-aligned int32 view
    m30 = P.ntt_primes(30, L)
    x = (rand_lbn(5, m30, B, N)).astype(np.int32)
    flat = torch.zeros(L * B * N + 4, dtype=torch.int32, device=cuda_device)
    v = flat[1:1 + L * B * N].view(L, B, N)
    v.copy_(torch.from_numpy(x).to(cuda_device))
    assert v.data_ptr() % 16 == 4
    assert np.array_equal(P.NTT(v, MOD=m30).cpu().numpy().astype(np.uint64), oracle_ring(N, m30).ntt(x.astype(np.uint64)))


def test_range_check_is_opt_in(cuda_device, monkeypatch):
    import polynomial as P
    from gpu_fhe_b200 import dispatch
    q = P.ntt_primes(60, 1)[0]
    x = torch.full((1, 1, 256), q, dtype=torch.int64, device=cuda_device)     # q itself: not canonical
    P.NTT(x, MOD=q)                                                           # default: precondition, unchecked
    monkeypatch.setattr(dispatch, "_CHECK_RANGE", True)
    with pytest.raises(ValueError, match="canonical"):
        P.NTT(x, MOD=q)
    with pytest.raises(ValueError, match="canonical"):
        P.poly_mul(x - 1, -x, q)
    ok = x - 1
    assert torch.equal(P.iNTT(P.NTT(ok, MOD=q), MOD=q), ok)


# ---------------------------------------------------------------------------------------------------
# serialisable plan cache
# ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("bits,N", [(60, 8192), (50, 4096), (30, 16384), (61, 256)])
def test_plan_save_load_roundtrip(cuda_device, tmp_path, bits, N):
    import primitive as P
    from gpu_fhe_b200 import _capi, ring
    L, B = 3, 2
    moduli = P.ntt_primes(bits, L)
    plan = P.make_plan(N, moduli, device=cuda_device.index)
    path = str(tmp_path / "plan.gfhe")
    P.save_plan(plan, path)
    assert os.path.getsize(path) == _capi.lib.gfhe_plan_export_bytes(plan.handle)
    loaded = P.load_plan(path, device=cuda_device.index)
    assert (loaded.N, loaded.moduli, loaded.psis) == (N, tuple(moduli), plan.psis)
    x, y = rand_lbn(4, moduli, B, N), rand_lbn(5, moduli, B, N)
    R = oracle_ring(N, moduli)
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    tx, ty = dev_i64(x, cuda_device), dev_i64(y, cuda_device)
    out = torch.empty_like(tx)
    assert _capi.lib.gfhe_ntt_forward(loaded.handle, 0, L, B, tx.data_ptr(), out.data_ptr(), st) == 0
    assert np.array_equal(host_u64(out), R.ntt(x))
    ws_bytes = _capi.lib.gfhe_poly_mul_workspace_bytes(loaded.handle, L, B)
    ws = torch.empty(max(ws_bytes, 16), dtype=torch.uint8, device=cuda_device)
    assert _capi.lib.gfhe_poly_mul(loaded.handle, 0, L, B, tx.data_ptr(), ty.data_ptr(), out.data_ptr(), ws.data_ptr(), ws_bytes, st) == 0
    assert np.array_equal(host_u64(out), R.poly_mul(x, y))
    loaded.close()
    # damaged images are refused
    raw = bytearray(open(path, "rb").read())
    raw[len(raw) // 2] ^= 0x40
    open(path, "wb").write(bytes(raw))
    with pytest.raises(_capi.GpuFheError) as ei:
        P.load_plan(path, device=cuda_device.index)
    assert ei.value.code == -14
    open(path, "wb").write(bytes(raw[:1000]))
    with pytest.raises(_capi.GpuFheError):
        P.load_plan(path, device=cuda_device.index)
    del ring


def test_plan_disk_cache(cuda_device, tmp_path, monkeypatch):
    import primitive as P
    from gpu_fhe_b200 import ring
    monkeypatch.setenv("GFHE_PLAN_CACHE_DIR", str(tmp_path))
    N, moduli = 4096, P.ntt_primes(60, 2)
    P.clear_plans()
    P.make_plan(N, moduli, device=cuda_device.index)
    files = os.listdir(tmp_path)
    assert len(files) == 1 and files[0].endswith(".gfhe")
    P.clear_plans()

    def no_search(*a, **k):
        raise AssertionError("root search ran although the plan image is cached on disk")
    monkeypatch.setattr(ring, "find_psi", no_search)
    x = rand_lbn(8, moduli, 2, N)
    got = P.NTT(dev_i64(x, cuda_device), MOD=moduli)           # plan comes from the image: no find_psi, no table generation
    assert np.array_equal(host_u64(got), oracle_ring(N, moduli).ntt(x))
    P.clear_plans()


# ---------------------------------------------------------------------------------------------------
# host-buffer entry points: pageable callers are staged, pinned ones are not; both pipelines are exact
# ---------------------------------------------------------------------------------------------------
def test_host_path_pageable_and_pinned(cuda_device):
    import primitive as P
    from gpu_fhe_b200 import dispatch
    N, L, B = 16384, 7, 36            # 4.5 MiB per limb: several groups per slot, ragged tail
    moduli = P.ntt_primes(60, L)
    a, b = rand_lbn(1, moduli, B, N), rand_lbn(2, moduli, B, N)
    want = oracle_ring(N, moduli).poly_mul(a, b, fast=True, nthreads=16)
    os.environ["GFHE_HOST_GROUP_MIB"] = "8"          # force many groups so slots are reused
    try:
        out = np.zeros_like(a)
        dispatch.poly_mul_host(a, b, out, moduli, device=cuda_device.index)
        st = dispatch.host_stats(moduli, N, device=cuda_device.index)
        assert np.array_equal(out, want)
        assert st["staged"] == 1.0 and st["groups"] >= 4 and st["h2d_bytes"] == 2 * a.nbytes and st["d2h_bytes"] == a.nbytes
        assert st["h2d_ms"] > 0 and st["d2h_ms"] > 0 and st["kernel_ms"] > 0 and st["stage_in_ms"] > 0
        pa = torch.from_numpy(a.view(np.int64)).pin_memory()
        pb = torch.from_numpy(b.view(np.int64)).pin_memory()
        pc = torch.zeros_like(pa).pin_memory()
        dispatch.poly_mul_host(pa.numpy().view(np.uint64), pb.numpy().view(np.uint64), pc.numpy().view(np.uint64), moduli,
                               device=cuda_device.index)
        st = dispatch.host_stats(moduli, N, device=cuda_device.index)
        assert np.array_equal(pc.numpy().view(np.uint64), want)
        assert st["staged"] == 0.0 and st["stage_in_ms"] == 0.0
        X = np.zeros_like(a)
        dispatch.ntt_host(a, X, moduli, device=cuda_device.index)
        back = np.zeros_like(a)
        dispatch.ntt_host(X, back, moduli, inverse=True, device=cuda_device.index)
        assert np.array_equal(back, a)
    finally:
        del os.environ["GFHE_HOST_GROUP_MIB"]


def test_host_path_groups_smaller_than_a_limb(cuda_device):
    """A group size below one limb's bytes cuts every limb's batch (small shards: more pipeline stages than limbs);
    uneven cuts (batch 5 in 3 parts), pinned and pageable callers, product and both transforms."""
    import primitive as P
    from gpu_fhe_b200 import dispatch
    N, L, B = 65536, 3, 5             # 2.5 MiB per limb, groups of 1 MiB: 3 parts of 1, 2, 2 polynomials per limb
    moduli = P.ntt_primes(60, L)
    a, b = rand_lbn(5, moduli, B, N), rand_lbn(6, moduli, B, N)
    R = oracle_ring(N, moduli)
    want = R.poly_mul(a, b, fast=True, nthreads=16)
    os.environ["GFHE_HOST_GROUP_MIB"] = "1"
    try:
        out = np.zeros_like(a)
        dispatch.poly_mul_host(a, b, out, moduli, device=cuda_device.index)
        st = dispatch.host_stats(moduli, N, device=cuda_device.index)
        assert np.array_equal(out, want)
        assert st["groups"] == 9 and st["staged"] == 1.0 and st["h2d_bytes"] == 2 * a.nbytes and st["d2h_bytes"] == a.nbytes
        pa = torch.from_numpy(a.view(np.int64)).pin_memory()
        pb = torch.from_numpy(b.view(np.int64)).pin_memory()
        pc = torch.zeros_like(pa).pin_memory()
        dispatch.poly_mul_host(pa.numpy().view(np.uint64), pb.numpy().view(np.uint64), pc.numpy().view(np.uint64), moduli,
                               device=cuda_device.index)
        st = dispatch.host_stats(moduli, N, device=cuda_device.index)
        assert np.array_equal(pc.numpy().view(np.uint64), want) and st["groups"] == 9 and st["staged"] == 0.0
        X = np.zeros_like(a)
        dispatch.ntt_host(a, X, moduli, device=cuda_device.index)
        assert np.array_equal(X, R.ntt(a, fast=True, nthreads=16))
        back = np.zeros_like(a)
        dispatch.ntt_host(X, back, moduli, inverse=True, device=cuda_device.index)
        assert np.array_equal(back, a)
    finally:
        del os.environ["GFHE_HOST_GROUP_MIB"]


# ---------------------------------------------------------------------------------------------------
# more than one GPU in the process / on the box
# ---------------------------------------------------------------------------------------------------
def test_plan_calls_keep_the_current_device(cuda_device):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import primitive as P
    N, moduli = 1024, P.ntt_primes(60, 2)
    torch.cuda.set_device(0)
    P.clear_plans()
    P.make_plan(N, moduli, device=1)
    assert torch.cuda.current_device() == 0
    x = rand_lbn(1, moduli, 2, N)
    t1 = dev_i64(x, torch.device("cuda", 1))
    got = P.NTT(t1, MOD=moduli)                    # tensors on device 1 while device 0 is current
    assert got.device.index == 1 and torch.cuda.current_device() == 0
    assert np.array_equal(host_u64(got), oracle_ring(N, moduli).ntt(x))
    P.clear_plans()
    assert torch.cuda.current_device() == 0


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _mgpu_worker(rank, world, port, L, B, N, bits, result_dir):
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import torch.distributed as dist

    import polynomial as P
    from gpu_fhe_b200 import shard
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    moduli = P.ntt_primes(bits, L)
    sh = shard.partition(L, B, world, rank)
    # every rank can regenerate any (limb, batch) block: one generator per limb
    def block(l, which):
        return np.random.default_rng([20261104, which, l]).integers(0, moduli[l], (B, N), dtype=np.uint64)
    mine = range(sh.limbs[0], sh.limbs[1])
    a = np.stack([block(l, 0)[sh.batch[0]:sh.batch[1]] for l in mine])
    b = np.stack([block(l, 1)[sh.batch[0]:sh.batch[1]] for l in mine])
    local = P.poly_mul(dev_i64(a, dev), dev_i64(b, dev), list(moduli[sh.limbs[0]:sh.limbs[1]]))
    full = shard.all_gather_limbs(local, sh)
    torch.cuda.synchronize()
    ok = tuple(full.shape) == (L, B, N)
    got = host_u64(full)
    # every rank checks one limb-polynomial from EVERY rank's shard; rank 0 checks the whole array
    for r in range(world):
        o = shard.partition(L, B, world, r)
        l, j = o.limbs[0] + (rank % o.num_limbs), o.batch[0] + (rank % o.num_batch)
        R1 = cref.Ring(N, [moduli[l]], [O.find_psi(moduli[l], N)])
        want = R1.poly_mul(block(l, 0)[None, j:j + 1], block(l, 1)[None, j:j + 1], fast=True)[0, 0]
        ok = ok and np.array_equal(got[l, j], want)
    if rank == 0:
        Rf = cref.Ring(N, moduli, [O.find_psi(q, N) for q in moduli])
        fa = np.stack([block(l, 0) for l in range(L)])
        fb = np.stack([block(l, 1) for l in range(L)])
        ok = ok and np.array_equal(got, Rf.poly_mul(fa, fb, fast=True, nthreads=8))
    open(os.path.join(result_dir, "rank%d" % rank), "w").write("ok" if ok else "bad")
    dist.destroy_process_group()


@pytest.mark.parametrize("L,B,N,bits", [(8, 4, 16384, 60), (2, 8, 4096, 50)])
def test_multi_gpu_sharded_poly_mul_all_gather(cuda_device, tmp_path, L, B, N, bits):
    """N-GPU parity on hardware: limb(+batch)-sharded poly_mul through the drop-in API on every visible GPU, NCCL
    all-gather, gathered [L, B, N] compared with the oracle on every rank (skips on a one-GPU box)."""
    world = min(torch.cuda.device_count(), 8)
    while world > 1 and (L * B) % world:
        world -= 1
    if world < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp
    mp.spawn(_mgpu_worker, args=(world, _free_port(), L, B, N, bits, str(tmp_path)), nprocs=world, join=True)
    for r in range(world):
        assert open(tmp_path / ("rank%d" % r)).read() == "ok", r


def _peer_worker(rank, world, port, L, B, N, bits, result_dir):
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import torch.distributed as dist

    import polynomial as P
    from gpu_fhe_b200 import shard
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    moduli = P.ntt_primes(bits, L)

    def block(l, which):
        return np.random.default_rng([77, which, l]).integers(0, moduli[l], (B, N), dtype=np.uint64)
    fa = np.stack([block(l, 0) for l in range(L)])
    fb = np.stack([block(l, 1) for l in range(L)])
    want = cref.Ring(N, moduli, [O.find_psi(q, N) for q in moduli]).poly_mul(fa, fb, fast=True, nthreads=4)
    ok = True
    for cyclic in (False, True):
        sh = shard.partition(L, B, world, rank, cyclic=cyclic)
        ids = sh.limb_ids
        a = np.stack([fa[l, sh.batch[0]:sh.batch[1]] for l in ids])
        b = np.stack([fb[l, sh.batch[0]:sh.batch[1]] for l in ids])
        mine = [moduli[l] for l in ids]
        pg = shard.PeerGather(L, B, N, dev)
        # (1) the whole block after the product
        local = P.poly_mul(dev_i64(a, dev), dev_i64(b, dev), mine)
        pg.full.zero_()
        torch.cuda.synchronize(); dist.barrier()
        pg.push(local, sh)
        pg.finish()
        torch.cuda.synchronize()
        ok = ok and np.array_equal(host_u64(pg.full), want)
        # (1b) the block produced in place inside this rank's own array: pushed to the peers only
        blk = pg.local_block(sh)
        ok = ok and ((blk is None) == (cyclic or sh.g_b != 1))
        if blk is not None:
            dist.barrier()
            pg.full.zero_()
            torch.cuda.synchronize(); dist.barrier()
            blk.copy_(local)
            pg.push(blk, sh)
            pg.finish()
            torch.cuda.synchronize()
            ok = ok and np.array_equal(host_u64(pg.full), want)
        # (2) limb by limb on a side stream while the next limb is multiplied
        dist.barrier()
        pg.full.zero_()
        torch.cuda.synchronize(); dist.barrier()
        side = torch.cuda.Stream(dev)
        outs = []
        for k in range(len(ids)):
            ck = P.poly_mul(dev_i64(a[k:k + 1], dev), dev_i64(b[k:k + 1], dev), [mine[k]])
            outs.append(ck)
            ev = torch.cuda.Event(); ev.record()
            side.wait_event(ev)
            with torch.cuda.stream(side):
                pg.push_limb(ck, 0, shard.Shard(sh.rank, sh.world, sh.g_l, sh.g_b, (ids[k], ids[k] + 1), sh.batch))
        ev = torch.cuda.Event(); ev.record(side)
        torch.cuda.current_stream().wait_event(ev)
        pg.finish()
        torch.cuda.synchronize()
        ok = ok and np.array_equal(host_u64(pg.full), want)
        pg.close()
    open(os.path.join(result_dir, "rank%d" % rank), "w").write("ok" if ok else "bad")
    dist.destroy_process_group()


@pytest.mark.parametrize("L,B,N,bits", [(8, 2, 8192, 60), (2, 8, 4096, 50)])
def test_multi_gpu_peer_store_gather(cuda_device, tmp_path, L, B, N, bits):
    """The library's own all-gather (NVLink peer stores, csrc/gather.cu through shard.PeerGather): block and cyclic limb maps,
    limb-and-batch shards, whole-block and limb-by-limb pushes; every rank's full array against the oracle."""
    world = min(torch.cuda.device_count(), 8)
    while world > 1 and (L * B) % world:
        world -= 1
    if world < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp
    mp.spawn(_peer_worker, args=(world, _free_port(), L, B, N, bits, str(tmp_path)), nprocs=world, join=True)
    for r in range(world):
        assert open(tmp_path / ("rank%d" % r)).read() == "ok", r
