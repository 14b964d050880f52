import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import polynomial as P
from oracle import oracle as O, cref
for N in (16, 256, 2048, 8192):
    q = P.ntt_primes(50, 1, N=4096)[0]; psi = O.find_psi(q, N)
    rng = np.random.default_rng(N)
    a = rng.integers(0, q, (1, 1, N), dtype=np.uint64); b = rng.integers(0, q, (1, 1, N), dtype=np.uint64)
    ta = torch.from_numpy(a.view(np.int64)).cuda(); tb = torch.from_numpy(b.view(np.int64)).cuda()
    R = cref.Ring(N, [q], [psi])
    with P.ring([q]):
        X = P.NTT(ta).cpu().numpy().view(np.uint64)
        want = R.ntt(a)
        bad = np.flatnonzero(X.ravel() != want.ravel())
        print("N", N, "NTT mismatches", bad.size, bad[:8], X.ravel()[:3], want.ravel()[:3])
        back = P.iNTT(torch.from_numpy(want.view(np.int64)).cuda()).cpu().numpy().view(np.uint64)
        bad = np.flatnonzero(back.ravel() != a.ravel())
        print("     iNTT(oracle NTT) mismatches", bad.size, bad[:8], back.ravel()[:3], a.ravel()[:3])
        c = P.poly_mul(ta, tb).cpu().numpy().view(np.uint64)
        wc = R.poly_mul(a, b)
        bad = np.flatnonzero(c.ravel() != wc.ravel())
        print("     polymul mismatches", bad.size, bad[:8])
