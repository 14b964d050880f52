#!/usr/bin/env python
"""One small shape per kernel family and arithmetic policy, for compute-sanitizer (racecheck / synccheck / memcheck).

    compute-sanitizer --tool racecheck python tools/sanitize.py > gpurun_out/sanitize_racecheck.log 2>&1

compute-sanitizer is CLOSED on this GPU pool (gpurun answers "closed on this pool and stays closed", profiles/
r02_sanitizer_closed.txt), so the substitute is `--stress N`: every case is run N times with identical inputs while a
second stream keeps the memory system and the SMs busy with unrelated copies (so CTA residency, warp interleaving and
memory latencies differ from run to run), and every run must reproduce the oracle's words bit for bit.  A missing barrier
or a mis-ordered mbarrier phase shows up as a run-to-run difference; none of the runs may differ.

The kernels elide some CTA barriers on no-hazard arguments (ntt_kernels.inl: the staged copy-out after the last forward
round; ntt_core.cuh: in-place exchanges) and hand-roll mbarrier phases for the TMA tile loads (k_row_polymul, the
persistent variant, the BNTT path): this is the tool's word on them.  Every result is also compared with the oracle, so a
run doubles as a parity check.  torch is used for allocation and copies only (no torch kernels run under the tool).
"""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402

import polynomial as P  # noqa: E402
from gpu_fhe_b200 import _capi  # noqa: E402
from oracle import cref  # noqa: E402
from oracle import oracle as O  # noqa: E402

lib = _capi.lib
dev = torch.device("cuda", 0)
st = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)


def to_dev(a, packed):
    h = np.ascontiguousarray(a).view(np.int64)
    if packed:
        h = h.astype(np.int32)
    return torch.from_numpy(h).to(dev)


def back(t, packed):
    h = t.cpu().numpy()
    return h.astype(np.uint64) if packed else h.view(np.uint64)


STRESS = 1
_noise = None


def noise():
    """unrelated work on a side stream: device-to-device copies of a few hundred MB, re-issued every time it is called"""
    global _noise
    if STRESS <= 1:
        return
    if _noise is None:
        _noise = (torch.cuda.Stream(dev), torch.empty(1 << 26, dtype=torch.int32, device=dev), torch.empty(1 << 26, dtype=torch.int32, device=dev))
    s, x, y = _noise
    with torch.cuda.stream(s):
        for _ in range(2):
            y.copy_(x, non_blocking=True)


def case(tag, logn, bits, L, B, packed=False, env=None):
    env = env or {}
    old = {k: os.environ.get(k) for k in env}
    os.environ.update(env)
    try:
        n = 1 << logn
        moduli = P.ntt_primes(bits, L)
        P.clear_plans()                                   # policy knobs are read at plan creation
        plan = P.make_plan(n, moduli, device=0)
        rng = np.random.default_rng(logn * 100 + bits)
        a = np.stack([rng.integers(0, q, (B, n), dtype=np.uint64) for q in moduli])
        b = np.stack([rng.integers(0, q, (B, n), dtype=np.uint64) for q in moduli])
        a[:, :, :4] = np.array(moduli, dtype=np.uint64).reshape(L, 1, 1) - 1
        R = cref.Ring(n, moduli, [O.find_psi(q, n) for q in moduli])
        ta, tb = to_dev(a, packed), to_dev(b, packed)
        tc, tx = torch.empty_like(ta), torch.empty_like(ta)
        f_fwd, f_inv, f_mul, f_pre = ((lib.gfhe_ntt_forward_u32, lib.gfhe_ntt_inverse_u32, lib.gfhe_poly_mul_u32, lib.gfhe_poly_mul_pre_u32)
                                      if packed else (lib.gfhe_ntt_forward, lib.gfhe_ntt_inverse, lib.gfhe_poly_mul, lib.gfhe_poly_mul_pre))
        wsb = lib.gfhe_poly_mul_workspace_bytes(plan.handle, L, B) // (2 if packed else 1)
        ws = torch.empty(max(wsb, 16), dtype=torch.uint8, device=dev)
        ok = [True] * 4
        X = R.ntt(a)
        want = R.poly_mul(a, b)
        tbn = to_dev(R.ntt(b), packed)
        for rep in range(STRESS):
            noise()
            assert f_fwd(plan.handle, 0, L, B, ta.data_ptr(), tx.data_ptr(), st) == 0
            ok[0] &= np.array_equal(back(tx, packed), X)
            noise()
            assert f_inv(plan.handle, 0, L, B, tx.data_ptr(), tc.data_ptr(), st) == 0
            ok[1] &= np.array_equal(back(tc, packed), a)
            noise()
            assert f_mul(plan.handle, 0, L, B, ta.data_ptr(), tb.data_ptr(), tc.data_ptr(), ws.data_ptr(), wsb, st) == 0
            ok[2] &= np.array_equal(back(tc, packed), want)
            noise()
            assert f_pre(plan.handle, 0, L, B, ta.data_ptr(), tbn.data_ptr(), tc.data_ptr(), st) == 0
            ok[3] &= np.array_equal(back(tc, packed), want)
        torch.cuda.synchronize()
        print("%-34s N=2^%-2d bits=%d L=%d B=%d  ntt/intt/polymul/polymul_pre: %s" % (tag, logn, bits, L, B, ok), flush=True)
        return all(ok)
    finally:
        for k, v in old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v


def main():
    global STRESS
    if "--stress" in sys.argv:
        STRESS = int(sys.argv[sys.argv.index("--stress") + 1])
    quick = "--quick" in sys.argv
    good = True
    policies = [("Lazy60", 60, False, {}), ("Harvey", 61, False, {}), ("Fp50", 50, False, {}), ("Shoup32", 30, False, {}),
                ("Shoup32P packed", 30, True, {})]
    sizes = [4, 5, 6, 7, 8, 10, 12, 13, 16, 17] if not quick else [5, 8, 12, 13, 16]
    for name, bits, packed, env in policies:
        for logn in sizes:
            good &= case(name, logn, bits, 2, 2 if logn <= 13 else 1, packed, env)
    # the persistent fused row pass (default for 32-bit words only) forced on for the 64-bit policies, and off for the 32-bit one
    for name, bits, packed, env in [("Lazy60 persist=1", 60, False, {"GFHE_PERSIST": "1"}), ("Harvey persist=1", 61, False, {"GFHE_PERSIST": "1"}),
                                    ("Fp50 persist=1", 50, False, {"GFHE_PERSIST": "1"}), ("Shoup32 persist=0", 30, False, {"GFHE_PERSIST": "0"}),
                                    ("Shoup32P persist=0", 30, True, {"GFHE_PERSIST": "0"})]:
        for logn in ([13, 16] if not quick else [16]):
            good &= case(name, logn, bits, 2, 3, packed, env)     # 6 polynomials: persistent CTAs loop over several tiles each
    # the pipelined single-launch kernels (in-kernel ticket / dependency waits)
    for name, bits, packed in [("Lazy60 pipe", 60, False), ("Fp50 pipe", 50, False), ("Shoup32 pipe", 30, False), ("Shoup32P pipe", 30, True)]:
        for logn, L, B in ([(13, 2, 40), (16, 3, 5), (17, 1, 3)] if not quick else [(16, 2, 3)]):
            good &= case(name, logn, bits, L, B, packed, {"GFHE_PIPE": "1"})
    # element-wise, multiply-accumulate, automorphism
    q = P.ntt_primes(60, 2)
    x = torch.from_numpy(np.arange(2 * 3 * 256, dtype=np.int64).reshape(2, 3, 256)).to(dev)
    s = P.vec_add(x, x, np.array(q, dtype=object).reshape(2, 1, 1))
    m = P.poly_mul_ntt(x, x, MOD=q)
    au = P.automorphism(x, 5, MOD=q)
    au2 = P.automorphism(x, 5, ntt=True, MOD=q)
    torch.cuda.synchronize()
    good &= bool((s.cpu() == (2 * x.cpu())).all()) and m.shape == x.shape and au.shape == au2.shape
    print("vec_add / mul_acc / automorphism ran", flush=True)
    print("SANITIZE_RESULT", "ok" if good else "MISMATCH", "(stress repetitions per case: %d)" % STRESS, flush=True)
    P.clear_plans()
    return 0 if good else 1


if __name__ == "__main__":
    sys.exit(main())
