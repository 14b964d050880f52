// Slab-level negacyclic NTT building blocks (device code).
//
// A "slab" is 2^LOGS <= 4096 residues staged on one CTA: each thread owns 16 residues in
// registers and runs up to four radix-2 stages ("a round") on them without touching memory;
// between rounds the slab is re-distributed through row-padded shared memory.
//
//   forward : Cooley-Tukey, natural in -> bit-reversed out, lazy residues in [0, 4q) (Harvey) / [0,16q) (Lazy60)
//   inverse : Gentleman-Sande, bit-reversed in -> natural out, lazy residues in [0, 2q) / [0, 8q),
//             N^-1 folded into the twiddles of the last stage
// The arithmetic policy A (modarith.cuh) is chosen per plan: Lazy60 when every modulus has four spare bits.
//
// The bit reversal is never materialised: it is absorbed by running the stages top bit first
// (forward) / bottom bit first (inverse) with bit-reversed twiddle tables.
//
// Semantics replace the identity stubs /root/reference/arithmetic.py:15-19; the contract is
// NTT(x)[k] = x(psi^(2*brv(k)+1)) mod q (oracle/oracle.py).
#pragma once
#include "modarith.cuh"

// timing experiments only (results are garbage when set): -DGFHE_EXP_NOSYNC drops the CTA barriers of the exchanges,
// -DGFHE_EXP_NOGLD replaces the global loads of residues by register arithmetic
#ifdef GFHE_EXP_NOSYNC
#define GFHE_SYNC() __syncwarp()
#else
#define GFHE_SYNC() __syncthreads()
#endif
#ifdef GFHE_EXP_NOGLD
#define GFHE_GLD(conv, expr, fake) conv((u64)(fake))
#else
#define GFHE_GLD(conv, expr, fake) conv(expr)
#endif

namespace gfhe {

constexpr int kElems = 16;       // residues per thread
constexpr int kMaxLogSlab = 12;  // 4096 residues = 32 KiB per slab

// Shared-memory placement of slab residue e: every 16-residue row (128 bytes) is followed by 16 bytes of
// padding, pad(e) = e + 2*(e/16) in 8-byte words.  A row-per-thread 16-byte access (window at bit 0) then
// spreads its 32 lanes over all banks 4 lanes deep (the minimum for 512 bytes), the 8-byte accesses of the
// other windows stay conflict-free -- and, unlike an XOR swizzle, pad() is ADDITIVE over disjoint bit fields:
// pad(thread bits | register bits) = pad(thread bits) + pad(register bits), so every register's slot is one
// per-thread base plus a compile-time immediate (no per-element address arithmetic, nothing to keep live).
__host__ __device__ constexpr u32 pad(u32 e) { return e + ((e >> 4) << 1); }
__host__ __device__ constexpr u32 padded_words(int logt) { return logt >= 4 ? pad(1u << logt) : (1u << logt); }

// Residue index inside the slab of register i of thread tid when the thread's 4-bit window
// sits at bit P0.
template <int P0>
__device__ __forceinline__ u32 elem_index(u32 tid, u32 i) {
    if constexpr (P0 == 0) return (tid << 4) | i;
    else return ((tid >> P0) << (P0 + 4)) | (i << P0) | (tid & ((1u << P0) - 1u));
}

// Read-only 16-byte twiddle load.  volatile: keeps the load inside the round that uses it -- a plain __ldg
// depends only on addresses, so the compiler hoists the loads of later transforms of a fused kernel to the
// top and then spills them.
__device__ __forceinline__ ulonglong2 ldg_tw(const ulonglong2* p) {
#ifdef GFHE_EXP_NOTW   // timing experiment only: no twiddle load (results are garbage)
    const u64 a = (u64)(uintptr_t)p;
    return make_ulonglong2(a | 1, a ^ 0x9E3779B97F4A7C15ull);
#endif
    ulonglong2 v;
    asm volatile("ld.global.nc.v2.u64 {%0, %1}, [%2];" : "=l"(v.x), "=l"(v.y) : "l"(p));
    return v;
}
// Compact 8-byte entry of the 32-bit policies: w | (w' << 32), w' = floor(w * 2^32 / q) = the high word of the 64-bit
// Shoup constant.  Returned in the positions Shoup32::mul reads ((u32)x and (u32)(y >> 32)): no instructions.
// The row passes read 240 bytes of 16-byte twiddles per thread against 128 bytes of data, all through L2 (tables of a
// 32-limb plan are 64 MB, re-read by every polynomial of the batch); with 32-bit words that is what bounds them.
__device__ __forceinline__ ulonglong2 tw_expand(ulonglong2 v) { return v; }
__device__ __forceinline__ ulonglong2 tw_expand(u64 e) { return make_ulonglong2(e & 0xffffffffull, e & 0xffffffff00000000ull); }
__device__ __forceinline__ ulonglong2 ldg_tw(const u64* p) {
    u64 e;
    asm volatile("ld.global.nc.u64 %0, [%1];" : "=l"(e) : "l"(p));
    return tw_expand(e);
}

// Where a round's twiddles come from.
//  Std : the bit-reversed table tw[k] = psi^brv(k); local bit b, sub-group hi reads tw[(v << (3-b)) + hi],
//        v = table index of the thread's group at the round's top stage.  Threads that share the bits above
//        the window share v, so a warp request touches 1-2 lines.
//  T0  : the window at bit 0 (each thread owns 16 contiguous residues and 15 private twiddles).  Reading
//        them from the standard table costs 32 lines per warp request; the plan therefore keeps a second,
//        warp-tiled copy  t0[t / 32][j][t % 32],  j = (2^(3-b) - 1) + hi in [0,15),  t = (residue index) / 16,
//        so the lanes of a warp read consecutive 16-byte entries and the 15 loads of a thread are ONE pointer
//        plus immediate offsets j * 512 bytes (no per-load address arithmetic, nothing to hoist and spill).
template <class E>
struct TwStd {
    const E* __restrict__ tw; u32 v;
    __device__ __forceinline__ ulonglong2 get(int b, int hi) const { return ldg_tw(tw + ((v << (3 - b)) + hi)); }
};
//  Sm   : the same indexing as Std, out of a copy of the table's first entries staged in shared memory (the
//        column kernels: all tiles of a limb use entries [1, 2^T) only, 4 KB).
template <class E>
struct TwSm {
    const E* s; u32 v;
    __device__ __forceinline__ ulonglong2 get(int b, int hi) const { return tw_expand(s[(v << (3 - b)) + hi]); }
};
template <class E>
struct TwT0 {
    const E* __restrict__ p;                  // t0 + (t / 32) * kT0Block + t % 32
    __device__ __forceinline__ ulonglong2 get(int b, int hi) const { return ldg_tw(p + ((((1 << (3 - b)) - 1) + hi) << 5)); }
};

// ---- one round of forward (CT) stages on local bits B_HI..B_LO --------------------------------
// v = (twiddle-table index of the thread's group at the round's top stage); the table index of
// local bit b, sub-group hi is (v << (3-b)) + hi  (see DESIGN.md "twiddle indexing").
// C = A::kFwdC * q.  Harvey: residues in [0,4q), X is brought under 2q every stage.  Lazy60: residues
// in [0,16q); the product is < 4q, so X grows by 4q per stage and is brought under 8q at the first
// and third stage of a round (every round therefore maps [0,16q) to [0,16q)).
// Lazy60 range schedule: the bound (in units of q) of every residue before stage `nst` of a chain of forward stages that
// starts from `entry`.  X is brought under 8q when the bound exceeds 12 (X + T must stay below 16q), then both outputs
// gain 4q.  From entry 16 this is "subtract at every second stage"; from canonical inputs (entry 1) the first three
// stages need nothing.
__host__ __device__ constexpr int fwd_bound(int entry, int nst) {
    int b = entry;
    for (int s = 0; s < nst; ++s) { if (b > 12) b = 8; b += 4; }
    return b;
}

template <class A, int B_HI, int B_LO, class TW, int ENTRY = 16>
__device__ __forceinline__ void fwd_round(typename A::word (&r)[kElems], const TW tw, typename A::word C, typename A::word nq) {
    typedef typename A::word W;
    if constexpr (A::kFloat) {
        // FP64 policy: C = q, nq = 1/q.  A product is within (0.5 + |y|/8q) q of zero (|y| < 2^51, q < 2^50).  X is brought
        // under q/2 at the first and third stage of a round only: with inputs below 1.9q the outputs of a reducing stage are
        // below 0.5 + 0.74 = 1.24q, those of the following stage below 1.24 + 0.5 + 1.24/8 = 1.9q < 2^51, so every round
        // maps (-1.9q, 1.9q) to itself.
#pragma unroll
        for (int sidx = 0; sidx <= B_HI - B_LO; ++sidx) {
            const int b = B_HI - sidx;
            const bool reduce_x = (sidx & 1) == 0;
#pragma unroll
            for (int hi = 0; hi < (8 >> b); ++hi) {
                const ulonglong2 t = tw.get(b, hi);
                const double w = bits_to_double(t.x), wq = bits_to_double(t.y);
#pragma unroll
                for (int lo = 0; lo < (1 << b); ++lo) {
                    const int ix = (hi << (b + 1)) + lo, iy = ix + (1 << b);
                    const double X = reduce_x ? fp_reduce(r[ix], C, nq) : r[ix];
                    const double T = fp_mulmod(r[iy], w, wq, C);
                    r[ix] = X + T;
                    r[iy] = X - T;
                }
            }
        }
        return;
    } else {
#pragma unroll
    for (int sidx = 0; sidx <= B_HI - B_LO; ++sidx) {
        const int b = B_HI - sidx;
        const bool reduce_x = !A::kLazy || fwd_bound(ENTRY, sidx) > 12;
#pragma unroll
        for (int hi = 0; hi < (8 >> b); ++hi) {
            const ulonglong2 t = tw.get(b, hi);
#pragma unroll
            for (int lo = 0; lo < (1 << b); ++lo) {
                const int ix = (hi << (b + 1)) + lo, iy = ix + (1 << b);
                const W X = reduce_x ? csub(r[ix], A::kLazy ? 2 * C : C) : r[ix];
                const W T = A::mul(r[iy], t.x, t.y, nq);
                r[ix] = X + T;
                r[iy] = X - T + C;
            }
        }
    }
    }
}

// Reduction of a forward residue that feeds the pointwise product: [0, 2q) is enough for Lazy60.
template <class A>
__device__ __forceinline__ typename A::word fwd_semi(typename A::word x, typename A::word q, typename A::word nq, u32 kshift, u32 kmu) {
    if constexpr (A::kFloat) return fp_reduce(x, q, nq);                   // |x| <= 0.51q
    else if constexpr (A::kLazy) return kred(x, kshift, kmu, nq);
    else return csub(csub(x, 2 * q), q);
}

// Full reduction of a forward residue to [0, q).
template <class A>
__device__ __forceinline__ typename A::word fwd_final(typename A::word x, typename A::word q, typename A::word nq, u32 kshift, u32 kmu) {
    if constexpr (A::kFloat) return fp_canon(fp_reduce(x, q, nq), q);     // -> [0, q)
    else if constexpr (A::kLazy) return csub(kred(x, kshift, kmu, nq), q);   // [0,16q) -> [0,2q) -> [0,q)
    else return csub(csub(x, 2 * q), q);                                // [0,4q)
}

// Static range tracking of the lazy inverse round: after a stage on local bit b-1, the registers with that bit set hold
// fresh products (< 4q) and the others sums (< 8q); the pair (ix, iy) of the next stage differs in bit b only, so both
// of its inputs are of the same kind.  A pair of products needs no conditional subtraction (12 of the 32 in a round).
__host__ __device__ constexpr bool inv_pair_is_fresh(int b, int b_lo, int lo) { return b > b_lo && ((lo >> (b - 1)) & 1) != 0; }

// The same idea for the FP64 policy.  Kind of register i BEFORE stage k of a round (stage k works on local bit b_lo + k):
//   0 = S, anything up to q - 1 in magnitude (round entry; an unreduced sum of two R),
//   1 = R, a reduced sum (<= (q-1)/2),   2 = P, a fresh product (<= 0.75q).
// Both inputs of a pair have the same kind.  R + R stays unreduced (it is an S); S + S and P + P are reduced (|sum| < 2q).
__host__ __device__ constexpr int fp_inv_kind(int k, int b_lo, int i) {
    int kind = 0;
    for (int j = 1; j <= k; ++j) kind = ((i >> (b_lo + j - 1)) & 1) ? 2 : (kind == 1 ? 0 : 1);
    return kind;
}

// ---- one round of inverse (GS) stages on local bits B_LO..B_HI --------------------------------
// C = A::kInvC * q: residues stay in [0, C) (Harvey 2q, Lazy60 8q).
// LAST: local bit B_HI is the final stage of the whole transform; it multiplies by N^-1
// (folded: U' = (U+V)*N^-1, V' = (U-V)*psi^-brv(1)*N^-1) with the exact product and fully reduces to [0, q).
// FRESH: the round's inputs are canonical residues (first round of a stand-alone inverse transform), so the first stage
// needs no conditional subtraction either.
template <class A, int B_HI, int B_LO, bool LAST, class TW, bool FRESH = false>
__device__ __forceinline__ void inv_round(typename A::word (&r)[kElems], const TW tw, typename A::word C, typename A::word nq,
                                          const LimbParams* __restrict__ Pg) {
    typedef typename A::word W;
    if constexpr (A::kFloat) {
        // FP64 policy: C = q, nq = 1/q.  Values entering a round are at most q - 1 in magnitude, so U - V and U + V stay
        // below 2q - 2 < 2^51; fp_inv_kind tracks which sums may stay unreduced (9 of the 32 in a round).
#pragma unroll
        for (int b = B_LO; b <= B_HI; ++b) {
            if (LAST && b == B_HI) {
                const double nd = __ldg(&Pg->ninv_d), nqd = __ldg(&Pg->ninv_q), wd = __ldg(&Pg->w1ninv_d), wqd = __ldg(&Pg->w1ninv_q);
#pragma unroll
                for (int lo = 0; lo < (1 << b); ++lo) {
                    const int ix = lo, iy = ix + (1 << b);
                    const double U = r[ix], V = r[iy];
                    r[ix] = fp_canon(fp_mulmod(U + V, nd, nqd, C), C);
                    r[iy] = fp_canon(fp_mulmod(U - V, wd, wqd, C), C);
                }
            } else {
#pragma unroll
                for (int hi = 0; hi < (8 >> b); ++hi) {
                    const ulonglong2 t = tw.get(b, hi);
                    const double w = bits_to_double(t.x), wq = bits_to_double(t.y);
#pragma unroll
                    for (int lo = 0; lo < (1 << b); ++lo) {
                        const int ix = (hi << (b + 1)) + lo, iy = ix + (1 << b);
                        const double U = r[ix], V = r[iy];
                        r[ix] = fp_inv_kind(b - B_LO, B_LO, ix) == 1 ? U + V : fp_reduce(U + V, C, nq);
                        r[iy] = fp_mulmod(U - V, w, wq, C);
                    }
                }
            }
        }
        return;
    } else {
#pragma unroll
    for (int b = B_LO; b <= B_HI; ++b) {
        if (LAST && b == B_HI) {
            // per-limb constants of the final stage are fetched here, not held in registers all along
            const ulonglong2 cn = __ldg(reinterpret_cast<const ulonglong2*>(&Pg->ninv));
            const ulonglong2 cw = __ldg(reinterpret_cast<const ulonglong2*>(&Pg->w1ninv));
            const W q = 0 - nq;
#pragma unroll
            for (int lo = 0; lo < (1 << b); ++lo) {                  // B_HI == 3 here: one group
                const int ix = lo, iy = ix + (1 << b);
                const W U = r[ix], V = r[iy];
                r[ix] = csub(A::mul_exact(U + V, cn.x, cn.y, nq), q);
                r[iy] = csub(A::mul_exact(U - V + C, cw.x, cw.y, nq), q);
            }
        } else {
#pragma unroll
            for (int hi = 0; hi < (8 >> b); ++hi) {
                const ulonglong2 t = tw.get(b, hi);
#pragma unroll
                for (int lo = 0; lo < (1 << b); ++lo) {
                    const int ix = (hi << (b + 1)) + lo, iy = ix + (1 << b);
                    const W U = r[ix], V = r[iy];
                    if (A::kLazy && (inv_pair_is_fresh(b, B_LO, lo) || (FRESH && b == B_LO))) {
                        // both inputs are products of the previous stage, < 4q: the sum is < 8q as it stands
                        r[ix] = U + V;
                        r[iy] = A::mul(U - V + (C >> 1), t.x, t.y, nq);
                    } else {
                        r[ix] = csub(U + V, C);                           // [0,C)
                        r[iy] = A::mul(U - V + C, t.x, t.y, nq);          // [0,2q) or [0,4q)
                    }
                }
            }
        }
    }
    }
}

// ---- shared-memory re-distribution ---------------------------------------------------------------
template <class W> struct __align__(16) W2 { W x, y; };   // 16-byte pair of 8-byte words (u64 or double)
// PK = padding words per 16-word row.  2 keeps 8-byte words (and the 8-byte-aligned pairs of 32-bit words the row kernels
// move) conflict-free; the column kernels' scalar exchange of 32-bit words wants 1 (two half-warps 272 = 16 mod 32 words apart).
template <int PK> __host__ __device__ constexpr u32 padk(u32 e) { return e + ((e >> 4) * PK); }
template <int P0, class W, int PK = 2>
__device__ __forceinline__ void sm_store(W* sm, u32 tid, const W (&r)[kElems]) {
    W* p = sm + padk<PK>(elem_index<P0>(tid, 0));
    if constexpr (P0 == 0 && sizeof(W) == 8) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            W2<W> v; v.x = r[2 * j]; v.y = r[2 * j + 1];
            *reinterpret_cast<W2<W>*>(p + 2 * j) = v;
        }
    } else {
#pragma unroll
        for (int i = 0; i < kElems; ++i) p[padk<PK>((u32)i << P0)] = r[i];
    }
}
template <int P0, class W, int PK = 2>
__device__ __forceinline__ void sm_load(const W* sm, u32 tid, W (&r)[kElems]) {
    const W* p = sm + padk<PK>(elem_index<P0>(tid, 0));
    if constexpr (P0 == 0 && sizeof(W) == 8) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const W2<W> v = *reinterpret_cast<const W2<W>*>(p + 2 * j);
            r[2 * j] = v.x; r[2 * j + 1] = v.y;
        }
    } else {
#pragma unroll
        for (int i = 0; i < kElems; ++i) r[i] = p[padk<PK>((u32)i << P0)];
    }
}

// ---- experiment (-DGFHE_EXP_SHFL): the north-star's "intra-warp butterflies with __shfl_sync" --------------------------
// The exchange between the window at bit 4 and the window at bit 0 is a 16 x 16 transpose inside every group of 16
// consecutive threads (register i of lane t <-> register t of lane i), so it CAN be done with warp shuffles instead of
// shared memory: four butterfly steps, each swapping half of the registers with lane ^ s.  Per thread that is 32
// exchanges x (2 selects + 1 shuffle + 2 selects) per 32-bit half -- ~8x the instructions of 16 stores + 16 loads, in
// kernels that are bound by instruction issue.  Kept as a measured A/B, not as the product path.
template <class W>
__device__ __forceinline__ W shfl_xor_word(W v, int s) {
    if constexpr (sizeof(W) == 8) {
        const unsigned long long b = sizeof(W) == 8 ? *reinterpret_cast<unsigned long long*>(&v) : 0ull;
        const unsigned long long o = __shfl_xor_sync(0xffffffffu, b, s);
        return *reinterpret_cast<const W*>(&o);
    } else {
        return __shfl_xor_sync(0xffffffffu, v, s);
    }
}
template <class W>
__device__ __forceinline__ void shfl_transpose16(W (&r)[kElems], u32 tid) {
#pragma unroll
    for (int s = 1; s < 16; s <<= 1) {
        const bool up = (tid & s) != 0;
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            if (i & s) continue;
            const W send = up ? r[i] : r[i | s];
            const W recv = shfl_xor_word(send, s);
            if (up) r[i] = recv; else r[i | s] = recv;
        }
    }
}

// Round schedule of a 2^LOGS contiguous slab: round k covers slab bits [p0(k), top(k)).
template <int LOGS>
struct Sched {
    static constexpr int NR = (LOGS + 3) / 4;
    static constexpr int top(int k) { return LOGS - 4 * k; }
    static constexpr int p0(int k) { return LOGS - 4 * (k + 1) > 0 ? LOGS - 4 * (k + 1) : 0; }
};

// Per-limb twiddle tables (already offset to the limb) and the ring degree.
template <class E>   // E = entry type: ulonglong2 {w, shoup(w)} / {w, w/q}, or u64 for the compact form of the 32-bit policies
struct TwTab {
    const E* __restrict__ tw;   // [N]       standard bit-reversed table
    const E* __restrict__ t0;   // [N/512][15][32] warp-tiled copy for the window at bit 0
    u32 N;
};
// Twiddle source of the round whose window sits at slab bit P0; off = offset of the slab in its polynomial.
template <int P0, class E>
__device__ __forceinline__ auto round_tw(const TwTab<E>& T, u32 off, u32 tid) {
    if constexpr (P0 == 0) {
        const u32 t = (off >> 4) + tid;
        return TwT0<E>{T.t0 + ((t >> 5) * kT0Block + (t & 31u))};
    }
    else return TwStd<E>{T.tw, ((T.N + off) >> (P0 + 4)) + (tid >> P0)};
}

// ---- forward transform of a slab: the LOGS low stages --------------------------------------------
// On entry r[] holds the layout of round 0 (window at bit p0(0)); on exit the layout of the last round
// (window at bit 0: 16 contiguous residues per thread).  The slab may be a tile of several independent
// 2^LOGS-point sub-transforms (tid then spans more than 2^(LOGS-4) threads): nothing here depends on it.
template <class A, int LOGS, int ENTRY = 16>
__device__ __forceinline__ void slab_fwd(typename A::word (&r)[kElems], typename A::word* sm, u32 tid, const TwTab<typename A::tw_t>& T, u32 off,
                                         typename A::word C, typename A::word nq) {
    using S = Sched<LOGS>;
    // stages done before round k: LOGS - top(k); ENTRY is the range bound of the slab's inputs (fwd_bound)
    {
        constexpr int P0 = S::p0(0);
        fwd_round<A, S::top(0) - 1 - P0, 0, decltype(round_tw<P0>(T, off, tid)), ENTRY>(r, round_tw<P0>(T, off, tid), C, nq);
    }
    if constexpr (S::NR > 1) {
        constexpr int PP = S::p0(0), P0 = S::p0(1);
#ifdef GFHE_EXP_SHFL
        if constexpr (PP == 4 && P0 == 0) shfl_transpose16(r, tid); else
#endif
        {
        sm_store<PP>(sm, tid, r);
        GFHE_SYNC();
        sm_load<P0>(sm, tid, r);
        }
        fwd_round<A, S::top(1) - 1 - P0, 0, decltype(round_tw<P0>(T, off, tid)), fwd_bound(ENTRY, LOGS - S::top(1))>(r, round_tw<P0>(T, off, tid), C, nq);
    }
    if constexpr (S::NR > 2) {
        constexpr int PP = S::p0(1), P0 = S::p0(2);
#ifdef GFHE_EXP_SHFL
        if constexpr (PP == 4 && P0 == 0) shfl_transpose16(r, tid); else
#endif
        {
        // no barrier needed: every thread overwrites exactly the words it read in the previous exchange
        sm_store<PP>(sm, tid, r);
        GFHE_SYNC();
        sm_load<P0>(sm, tid, r);
        }
        fwd_round<A, S::top(2) - 1 - P0, 0, decltype(round_tw<P0>(T, off, tid)), fwd_bound(ENTRY, LOGS - S::top(2))>(r, round_tw<P0>(T, off, tid), C, nq);
    }
}

// ---- inverse transform of a slab: the LOGS low stages, bottom bit first --------------------------
// Entry layout: window at bit 0; exit layout: window at bit p0(0).
// LAST: the slab is the whole polynomial (N == 2^LOGS), so the top stage folds N^-1.
template <class A, int LOGS, bool LAST, bool FRESH = false>
__device__ __forceinline__ void slab_inv(typename A::word (&r)[kElems], typename A::word* sm, u32 tid, const TwTab<typename A::tw_t>& T, u32 off,
                                         typename A::word C, typename A::word nq, const LimbParams* __restrict__ Pg) {
    using S = Sched<LOGS>;
    if constexpr (S::NR > 2) {
        constexpr int P0 = S::p0(2), PN = S::p0(1);
        inv_round<A, S::top(2) - 1 - P0, 0, false, decltype(round_tw<P0>(T, off, tid)), FRESH>(r, round_tw<P0>(T, off, tid), C, nq, Pg);
        sm_store<P0>(sm, tid, r);
        GFHE_SYNC();
        sm_load<PN>(sm, tid, r);
    }
    if constexpr (S::NR > 1) {
        constexpr int P0 = S::p0(1), PN = S::p0(0);
        inv_round<A, S::top(1) - 1 - P0, 0, false, decltype(round_tw<P0>(T, off, tid)), FRESH && S::NR == 2>(r, round_tw<P0>(T, off, tid), C, nq, Pg);
        sm_store<P0>(sm, tid, r);
        GFHE_SYNC();
        sm_load<PN>(sm, tid, r);
    }
    {
        constexpr int P0 = S::p0(0);
        inv_round<A, S::top(0) - 1 - P0, 0, LAST, decltype(round_tw<P0>(T, off, tid)), FRESH && S::NR == 1>(r, round_tw<P0>(T, off, tid), C, nq, Pg);
    }
}

}  // namespace gfhe
