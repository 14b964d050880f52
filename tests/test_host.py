"""CPU suite, part 3: host-side logic of the drop-in modules (no GPU)."""
import inspect

import numpy as np
import pytest


def test_drop_in_surface_matches_reference(built_lib):
    # the seven reference callables with their exact positional signatures
    import arithmetic
    import polynomial
    import primitive
    want = {"vec_add": ["a", "b", "MOD"], "vec_sub": ["a", "b", "MOD"], "vec_mul": ["a", "b", "MOD"],
            "NTT": ["x"], "iNTT": ["x"], "XXX": [], "poly_add": ["a", "b", "MOD"]}
    for name, params in want.items():
        fn = getattr(polynomial, name)
        positional = [p.name for p in inspect.signature(fn).parameters.values()
                      if p.kind in (p.POSITIONAL_ONLY, p.POSITIONAL_OR_KEYWORD)]
        assert positional == params, name
    # star-import chain: every lower name is exported by every module above it (incl. the leaked np)
    ns = {}
    exec("from polynomial import *", ns)
    for name in list(want) + ["np", "poly_mul"]:
        assert name in ns
    ns = {}
    exec("from primitive import *", ns)
    assert "poly_add" not in ns and "XXX" in ns and "NTT" in ns
    assert arithmetic.np is np and primitive.vec_add is arithmetic.vec_add


def test_xxx_prints_xxx(built_lib, capsys):
    import primitive
    assert primitive.XXX() is None
    assert capsys.readouterr().out == "XXX\n"       # /root/reference/primitive.py:3-4


# (q, N, smallest base x, psi) verified there with sympy -- literals, so that neither the product
# code's root search nor the oracle's can vouch for the other
ROOT_LITERALS = [
    (1073479681, 1024, 11, 530309095),
    (1125899902124033, 4096, 3, 93790016731446),
    (1152921504606584833, 65536, 5, 987813353222176621),
    (1152921504606584833, 131072, 5, 720645352895426071),
]


@pytest.mark.parametrize("q,N,x,psi", ROOT_LITERALS)
def test_root_literals_and_exact_order(built_lib, q, N, x, psi):
    import primitive as P
    from oracle import oracle as O
    assert P.find_psi(q, N) == psi                    # product code against the literal ...
    assert O.find_psi(q, N) == psi                    # ... and the oracle against the literal, each on its own
    assert psi == pow(x, (q - 1) // (2 * N), q)
    for y in range(2, x):                             # x is the SMALLEST base that works
        assert pow(pow(y, (q - 1) // (2 * N), q), N, q) != q - 1
    # psi has order exactly 2N: psi^(2N) = 1 and, 2N being a power of two, psi^N != 1 rules out every proper divisor
    assert pow(psi, 2 * N, q) == 1 and pow(psi, N, q) == q - 1 != 1
    assert len({pow(psi, k, q) for k in range(0, 2 * N, max(1, N // 64))}) == len(range(0, 2 * N, max(1, N // 64)))


def test_roots_and_primes(built_lib):
    import primitive as P
    # SURVEY.md section 8d prime families (literals from there)
    assert P.ntt_primes(60, 4) == [1152921504606584833, 1152921504598720513, 1152921504592429057, 1152921504581419009]
    assert P.ntt_primes(60, 64)[-1] == 1152921504241942529
    assert P.ntt_primes(50, 4) == [1125899902124033, 1125899886395393, 1125899884036097, 1125899882987521]
    assert P.ntt_primes(30, 2) == [1073479681, 1068236801]
    assert P.ntt_primes(30, 1, N=1024) == [1073707009]    # largest 30-bit prime = 1 mod 2048
    with pytest.raises(ValueError):
        P.find_psi(101, 16)
    assert P.is_probable_prime(1152921504606584833) and not P.is_probable_prime(1152921504606584833 - 4)


def test_ring_context(built_lib):
    import primitive as P
    from gpu_fhe_b200 import ring
    assert P.get_ring() is None
    with pytest.raises(RuntimeError):
        ring.resolve_moduli(None)                      # NTT(x) with no ring: clear error, no identity fallback
    with P.ring([97, 193]):
        assert ring.resolve_moduli(None) == (97, 193)
        with P.ring(np.array([[257]])):
            assert P.get_ring() == (257,)
        assert P.get_ring() == (97, 193)
    assert P.get_ring() is None
    assert ring.resolve_moduli(np.array([5, 7], dtype=object).reshape(2, 1, 1)) == (5, 7)
    assert ring.resolve_moduli(12289) == (12289,)


def test_shape_and_modulus_mapping(built_lib):
    import arithmetic  # noqa: F401
    from gpu_fhe_b200 import dispatch as d
    assert d._lbn((4, 16, 64), 4) == (4, 16, 64)
    assert d._lbn((16, 64), 1) == (1, 16, 64)
    assert d._lbn((64,), 1) == (1, 1, 64)
    assert d._lbn((2, 3, 5, 64), 2) == (2, 15, 64)
    with pytest.raises(ValueError):
        d._lbn((3, 64), 2)
    # MOD follows numpy's broadcasting rules (/root/reference/arithmetic.py:5): (moduli, result shape, axis permutation, inner)
    assert d._vec_moduli(7, (2, 3, 4)) == ((7,), (2, 3, 4), None, 24)
    assert d._vec_moduli(np.array([[[5]], [[7]]], dtype=object), (2, 3, 4)) == ((5, 7), (2, 3, 4), None, 12)
    assert d._vec_moduli(np.array([5, 7]), (2,)) == ((5, 7), (2,), None, 1)
    assert d._vec_moduli(np.array([5, 7]), (2, 3, 2)) == ((5, 7), (2, 3, 2), [2, 0, 1], 6)          # varies along the LAST axis
    assert d._vec_moduli(np.array([[5], [7], [11]]), (2, 3, 2)) == ((5, 7, 11), (2, 3, 2), [1, 0, 2], 4)
    assert d._vec_moduli(np.array([[[5]], [[7]]]), (3, 4)) == ((5, 7), (2, 3, 4), None, 12)        # MOD adds an axis
    assert d._vec_moduli(np.array([[5, 7], [11, 13]]), (2, 2)) == ((5, 7, 11, 13), (2, 2), None, 1)  # one modulus per element
    with pytest.raises(ValueError):          # does not broadcast: numpy raises ValueError too
        d._vec_moduli(np.array([5, 7, 11]), (2, 3, 2))


def test_no_cpu_fallback(built_lib):
    import torch

    import polynomial as P
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    a = np.arange(16, dtype=np.int64)
    for call in (lambda: P.vec_add(a, a, 97), lambda: P.NTT(a, MOD=97), lambda: P.poly_mul(a, a, 97)):
        with pytest.raises(RuntimeError, match="no CPU fallback"):
            call()
    with pytest.raises(AssertionError):     # shape mismatch: bare AssertionError like arithmetic.py:4
        P.vec_add(np.arange(4), np.arange(5), 7)


def test_object_operands_must_fit_a_word(built_lib):
    import arithmetic  # noqa: F401
    from gpu_fhe_b200 import dispatch as d
    # exact Python ints are accepted only if they fit ONE 64-bit reading (signed or unsigned); validated before any device work
    for bad in ([1 << 64], [-(1 << 63) - 1], [-1, 1 << 63]):
        with pytest.raises(ValueError, match="64-bit word"):
            d._Arg(np.array(bad, dtype=object))
    with pytest.raises(TypeError):
        d._Arg(np.array([0.5, 1.5]))


def test_host_entry_points_validate_shapes(built_lib):
    import arithmetic  # noqa: F401
    from gpu_fhe_b200 import dispatch as d
    a = np.zeros((2, 3, 16), dtype=np.uint64)
    # the C side copies L*B*N words from / to each array: mismatched shapes must never reach it
    with pytest.raises(AssertionError):
        d.poly_mul_host(a, np.ascontiguousarray(a[:, :2]), a.copy(), [97, 193])
    with pytest.raises(AssertionError):
        d.poly_mul_host(a, a, np.zeros((2, 3, 8), dtype=np.uint64), [97, 193])
    with pytest.raises(AssertionError):
        d.ntt_host(a, np.zeros((2, 2, 16), dtype=np.uint64), [97, 193])
    with pytest.raises(ValueError):                      # leading dimension is not the limb count
        d.poly_mul_host(a, a, a.copy(), [97, 193, 257])
    with pytest.raises(TypeError):
        d.poly_mul_host(a.astype(np.int32), a.astype(np.int32), a.astype(np.int32), [97, 193])
    ro = a.copy()
    ro.flags.writeable = False
    with pytest.raises(ValueError, match="writable"):
        d.poly_mul_host(a, a, ro, [97, 193])


def test_partition():
    import arithmetic  # noqa: F401
    from gpu_fhe_b200 import shard
    # config 4 on 8 GPUs: 4 limbs x 16 batch each;  config 3 on 2 GPUs: 4 limbs each
    s = [shard.partition(32, 16, 8, r) for r in range(8)]
    assert all(x.g_l == 8 and x.g_b == 1 and x.num_limbs == 4 and x.num_batch == 16 for x in s)
    assert [x.limbs for x in s] == [(4 * r, 4 * r + 4) for r in range(8)]
    s = [shard.partition(8, 128, 2, r) for r in range(2)]
    assert [x.limbs for x in s] == [(0, 4), (4, 8)]
    # fewer limbs than ranks: split the batch too
    s = [shard.partition(2, 8, 8, r) for r in range(8)]
    assert all(x.g_l == 2 and x.g_b == 4 for x in s)
    cover = set()
    for x in s:
        for l in range(*x.limbs):
            for b in range(*x.batch):
                assert (l, b) not in cover
                cover.add((l, b))
    assert len(cover) == 16
    with pytest.raises(ValueError):
        shard.partition(1, 3, 2, 0)
    # cyclic limb map: rank i holds limbs i, i+g, ...; the k-th local limbs of all ranks are the slab [k*g, (k+1)*g)
    s = [shard.partition(32, 16, 8, r, cyclic=True) for r in range(8)]
    assert [x.limb_ids for x in s][:2] == [(0, 8, 16, 24), (1, 9, 17, 25)] and all(x.num_limbs == 4 for x in s)
    for k in range(4):
        assert sorted(x.limb_ids[k] for x in s) == list(range(8 * k, 8 * k + 8))
    s = [shard.partition(2, 8, 8, r, cyclic=True) for r in range(8)]
    assert sorted((x.limb_ids, x.batch) for x in s) == sorted(((i,), (2 * j, 2 * j + 2)) for i in range(2) for j in range(4))


def test_device_map_exchange_between_concurrent_processes(built_lib, tmp_path):
    """Four processes started together, the way a launcher starts ranks (local rank 0 last, and it may finish before the
    others have looked): all four must come back with the same map."""
    import json
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = (
        "import sys, json; sys.path.insert(0, %r)\n"
        "import arithmetic\n"
        "from gpu_fhe_b200 import shard\n"
        "shard.torch.cuda.device_count = lambda: 8\n"
        "r = int(sys.argv[1])\n"
        "dev, res = shard.choose_local_devices(4, r, timeout_s=30)\n"
        "print(json.dumps([r, dev, res['devices']]))\n" % root)
    env = dict(os.environ, GFHE_DEVICE_MAP="probe", MASTER_PORT="29873", TMPDIR=str(tmp_path))
    procs = [subprocess.Popen([sys.executable, "-c", code, str(r)], env=env, stdout=subprocess.PIPE, text=True) for r in (3, 2, 1, 0)]
    outs = [json.loads(p.communicate(timeout=120)[0].strip().splitlines()[-1]) for p in procs]
    assert all(p.returncode == 0 for p in procs)
    maps = {tuple(o[2]) for o in outs}
    assert len(maps) == 1 and all(o[1] == o[2][o[0]] for o in outs)
    assert sorted(o[1] for o in outs) == list(maps.pop())


def test_numa_binding_helper_is_harmless_without_nvml(built_lib):
    """shard.bind_to_gpu_numa returns None (and changes nothing) when NVML has no device to ask."""
    import os
    import arithmetic  # noqa: F401
    from gpu_fhe_b200 import shard
    before = os.sched_getaffinity(0)
    prev = shard.bind_to_gpu_numa(0)
    if prev is None:
        assert os.sched_getaffinity(0) == before
    else:
        os.sched_setaffinity(0, prev)


def test_device_map_is_identity_unless_asked_and_ranks_agree(built_lib, monkeypatch):
    """shard.choose_local_devices: device = local rank by default and on a box without spare GPUs; with
    GFHE_DEVICE_MAP=probe local rank 0 always leaves an answer (the identity map when the probe cannot run, as here) and the
    other ranks read the same one."""
    import arithmetic  # noqa: F401
    from gpu_fhe_b200 import shard
    monkeypatch.delenv("GFHE_DEVICE_MAP", raising=False)
    assert shard.choose_local_devices(4, 2) == (2, None)
    monkeypatch.setenv("GFHE_DEVICE_MAP", "probe")
    assert shard.choose_local_devices(1, 0) == (0, None)
    monkeypatch.setattr(shard.torch.cuda, "device_count", lambda: 4)
    assert shard.choose_local_devices(4, 3) == (3, None)          # no second block of 4
    monkeypatch.setattr(shard.torch.cuda, "device_count", lambda: 8)
    monkeypatch.setenv("MASTER_PORT", "29871")
    dev0, res0 = shard.choose_local_devices(4, 0, timeout_s=60)
    assert res0["devices"] == [0, 1, 2, 3] or len(res0["devices"]) == 4
    for r in (1, 2, 3):
        dev, res = shard.choose_local_devices(4, r, timeout_s=1)
        assert res["devices"] == res0["devices"] and dev == res0["devices"][r]
    with pytest.raises(RuntimeError):                               # another launcher port: nobody wrote a map
        monkeypatch.setenv("MASTER_PORT", "29872")
        shard.choose_local_devices(4, 1, timeout_s=-59.5)
    import os
    import tempfile
    os.remove(os.path.join(tempfile.gettempdir(), "gfhe_devmap_29871.json"))
