"""Python model of the CUDA kernels' index math (ntt_core.cuh / ntt_kernels.cu): round schedule,
thread<->residue mapping, twiddle-table indexing, column geometry.  Runs the same data movement
with exact Python-int arithmetic and compares against the oracle, so indexing bugs are found on
the CPU before a GPU call is spent.  Dev tool; not part of the product path."""
import random, sys, os
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
from oracle import oracle as O

def brv(k, bits):
    return int(format(k, "0%db" % bits)[::-1], 2) if bits else 0

def tables(q, logN, psi):
    N = 1 << logN
    pinv = pow(psi, -1, q)
    fwd = [pow(psi, brv(k, logN), q) for k in range(N)]
    inv = [pow(pinv, brv(k, logN), q) for k in range(N)]
    return fwd, inv

def elem_index(P0, tid, i):
    if P0 == 0: return (tid << 4) | i
    return ((tid >> P0) << (P0 + 4)) | (i << P0) | (tid & ((1 << P0) - 1))

def fwd_round(r, tw, v, q, B_HI, B_LO):
    for b in range(B_HI, B_LO - 1, -1):
        for hi in range(8 >> b):
            w = tw[(v << (3 - b)) + hi]
            for lo in range(1 << b):
                ix = (hi << (b + 1)) + lo; iy = ix + (1 << b)
                X, T = r[ix], r[iy] * w % q
                r[ix], r[iy] = (X + T) % q, (X - T) % q

def inv_round(r, tw, v, q, B_HI, B_LO, last, ninv, w1ninv):
    for b in range(B_LO, B_HI + 1):
        if last and b == B_HI:
            for lo in range(1 << b):
                ix = lo; iy = ix + (1 << b)
                U, V = r[ix], r[iy]
                r[ix], r[iy] = (U + V) * ninv % q, (U - V) * w1ninv % q
        else:
            for hi in range(8 >> b):
                w = tw[(v << (3 - b)) + hi]
                for lo in range(1 << b):
                    ix = (hi << (b + 1)) + lo; iy = ix + (1 << b)
                    U, V = r[ix], r[iy]
                    r[ix], r[iy] = (U + V) % q, (U - V) * w % q

def sched(LOGS):
    NR = (LOGS + 3) // 4
    return [(max(LOGS - 4 * (k + 1), 0), LOGS - 4 * k) for k in range(NR)]   # (p0, top)

def slab_fwd(slab, LOGS, tw, G, q):
    NT = max(1 << (LOGS - 4), 1)
    for (P0, top) in sched(LOGS):
        for tid in range(NT):
            idx = [elem_index(P0, tid, i) for i in range(16)]
            r = [slab[e] for e in idx]
            fwd_round(r, tw, (G >> (P0 + 4)) + (tid >> P0), q, top - 1 - P0, 0)
            for e, val in zip(idx, r): slab[e] = val

def slab_inv(slab, LOGS, tw, G, q, last, ninv, w1ninv):
    NT = max(1 << (LOGS - 4), 1)
    S = sched(LOGS)
    for k in range(len(S) - 1, -1, -1):
        P0, top = S[k]
        for tid in range(NT):
            idx = [elem_index(P0, tid, i) for i in range(16)]
            r = [slab[e] for e in idx]
            inv_round(r, tw, (G >> (P0 + 4)) + (tid >> P0), q, top - 1 - P0, 0, last and k == 0, ninv, w1ninv)
            for e, val in zip(idx, r): slab[e] = val

def col_addr(LOGC, base, e):
    return base + (e & ((1 << LOGC) - 1)) + ((e >> LOGC) << 12)

def col_fwd(x, logN, tw, q):
    LOGC = 24 - logN
    for tile in range(1 << (12 - LOGC)):
        base = tile << LOGC
        windows = [(8, max(LOGC - 8, 0), lambda tid: 1)]
        if LOGC < 8: windows.append((4, max(LOGC - 4, 0), lambda tid: 16 + (tid >> 4)))
        for P0, lo_b, vf in windows:
            for tid in range(256):
                idx = [col_addr(LOGC, base, elem_index(P0, tid, i)) for i in range(16)]
                r = [x[a] for a in idx]
                fwd_round(r, tw, vf(tid), q, 3, lo_b)
                for a, val in zip(idx, r): x[a] = val

def col_inv(x, logN, tw, q, ninv, w1ninv):
    LOGC = 24 - logN
    for tile in range(1 << (12 - LOGC)):
        base = tile << LOGC
        windows = []
        if LOGC < 8: windows.append((4, max(LOGC - 4, 0), lambda tid: 16 + (tid >> 4), False))
        windows.append((8, max(LOGC - 8, 0), lambda tid: 1, True))
        for P0, lo_b, vf, last in windows:
            for tid in range(256):
                idx = [col_addr(LOGC, base, elem_index(P0, tid, i)) for i in range(16)]
                r = [x[a] for a in idx]
                inv_round(r, tw, vf(tid), q, 3, lo_b, last, ninv, w1ninv)
                for a, val in zip(idx, r): x[a] = val

def ntt_fwd(x, logN, q, psi):
    N = 1 << logN; fwd, _ = tables(q, logN, psi); x = list(x)
    if logN <= 12:
        slab_fwd(x, logN, fwd, N, q)
    else:
        col_fwd(x, logN, fwd, q)
        for row in range(N >> 12):
            s = x[row << 12:(row + 1) << 12]; slab_fwd(s, 12, fwd, N + (row << 12), q); x[row << 12:(row + 1) << 12] = s
    return x

def ntt_inv(x, logN, q, psi):
    N = 1 << logN; _, inv = tables(q, logN, psi); x = list(x)
    ninv = pow(N, -1, q); w1ninv = inv[1] * ninv % q
    if logN <= 12:
        slab_inv(x, logN, inv, N, q, True, ninv, w1ninv)
    else:
        for row in range(N >> 12):
            s = x[row << 12:(row + 1) << 12]; slab_inv(s, 12, inv, N + (row << 12), q, False, ninv, w1ninv); x[row << 12:(row + 1) << 12] = s
        col_inv(x, logN, inv, q, ninv, w1ninv)
    return x

if __name__ == "__main__":
    rng = random.Random(7)
    sizes = [int(a) for a in sys.argv[1:]] or [4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16]
    for logN in sizes:
        N = 1 << logN
        q = O.ntt_primes(60, 1, max(18, logN + 1))[0] if logN > 10 else O.ntt_primes(30, 1, 18)[0]
        psi = O.find_psi(q, N)
        x = O.random_poly(rng, N, q)
        X = ntt_fwd(x, logN, q, psi)
        ref = O.ntt_fast(x, q, psi)
        assert X == ref, "fwd mismatch logN=%d" % logN
        assert ntt_inv(X, logN, q, psi) == x, "inv mismatch logN=%d" % logN
        print("emulated schedule ok: logN =", logN)
