// Slab-level negacyclic NTT building blocks (device code).
//
// A "slab" is 2^LOGS <= 4096 residues staged on one CTA: each thread owns 16 residues in
// registers and runs up to four radix-2 stages ("a round") on them without touching memory;
// between rounds the slab is re-distributed through XOR-swizzled shared memory.
//
//   forward : Cooley-Tukey, natural in -> bit-reversed out, lazy residues in [0, 4q)
//   inverse : Gentleman-Sande, bit-reversed in -> natural out, lazy residues in [0, 2q),
//             N^-1 folded into the twiddles of the last stage
//
// The bit reversal is never materialised: it is absorbed by running the stages top bit first
// (forward) / bottom bit first (inverse) with bit-reversed twiddle tables.
//
// Semantics replace the identity stubs /root/reference/arithmetic.py:15-19; the contract is
// NTT(x)[k] = x(psi^(2*brv(k)+1)) mod q (oracle/oracle.py).
#pragma once
#include "modarith.cuh"

namespace gfhe {

constexpr int kElems = 16;       // residues per thread
constexpr int kMaxLogSlab = 12;  // 4096 residues = 32 KiB per slab

// 128-byte XOR swizzle at 16-byte-chunk granularity (same pattern as TMA SWIZZLE_128B):
// chunk bits [1,4) of the residue index are XORed with row bits [4,7).
__device__ __forceinline__ u32 swz(u32 e) { return e ^ (((e >> 4) & 7u) << 1); }

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
    ulonglong2 v;
    asm volatile("ld.global.nc.v2.u64 {%0, %1}, [%2];" : "=l"(v.x), "=l"(v.y) : "l"(p));
    return v;
}

// Where a round's twiddles come from.
//  Std : the bit-reversed table tw[k] = psi^brv(k); local bit b, sub-group hi reads tw[(v << (3-b)) + hi],
//        v = table index of the thread's group at the round's top stage.  Threads that share the bits above
//        the window share v, so a warp request touches 1-2 lines.
//  T0  : the window at bit 0 (each thread owns 16 contiguous residues and 15 private twiddles).  Reading
//        them from the standard table costs 32 lines per warp request; the plan therefore keeps a second,
//        transposed copy  t0[j][t],  j = (2^(3-b) - 1) + hi in [0,15),  t = (residue index) / 16,  so the
//        lanes of a warp read consecutive 16-byte entries.
struct TwStd {
    const ulonglong2* __restrict__ tw; u32 v;
    __device__ __forceinline__ ulonglong2 get(int b, int hi) const { return ldg_tw(tw + ((v << (3 - b)) + hi)); }
};
struct TwT0 {
    const ulonglong2* __restrict__ p; u32 stride;     // p = t0 + t, stride = N/16
    __device__ __forceinline__ ulonglong2 get(int b, int hi) const { return ldg_tw(p + (size_t)(((1 << (3 - b)) - 1) + hi) * stride); }
};

// ---- one round of forward (CT) stages on local bits B_HI..B_LO --------------------------------
// v = (twiddle-table index of the thread's group at the round's top stage); the table index of
// local bit b, sub-group hi is (v << (3-b)) + hi  (see DESIGN.md "twiddle indexing").
template <int B_HI, int B_LO, class TW>
__device__ __forceinline__ void fwd_round(u64 (&r)[kElems], const TW tw, u64 q2, u64 nq) {
#pragma unroll
    for (int sidx = 0; sidx <= B_HI - B_LO; ++sidx) {
        const int b = B_HI - sidx;
#pragma unroll
        for (int hi = 0; hi < (8 >> b); ++hi) {
            const ulonglong2 t = tw.get(b, hi);
#pragma unroll
            for (int lo = 0; lo < (1 << b); ++lo) {
                const int ix = (hi << (b + 1)) + lo, iy = ix + (1 << b);
                const u64 X = csub(r[ix], q2);                       // [0,4q) -> [0,2q)
                const u64 T = shoup_mul(r[iy], t.x, t.y, nq);       // [0,2q)
                r[ix] = X + T;                                        // [0,4q)
                r[iy] = X - T + q2;                                   // [0,4q)
            }
        }
    }
}

// ---- one round of inverse (GS) stages on local bits B_LO..B_HI --------------------------------
// LAST: local bit B_HI is the final stage of the whole transform; it multiplies by N^-1
// (folded: U' = (U+V)*N^-1, V' = (U-V)*psi^-brv(1)*N^-1) and fully reduces to [0, q).
template <int B_HI, int B_LO, bool LAST, class TW>
__device__ __forceinline__ void inv_round(u64 (&r)[kElems], const TW tw, u64 q2, u64 nq, const LimbParams* __restrict__ Pg) {
#pragma unroll
    for (int b = B_LO; b <= B_HI; ++b) {
        if (LAST && b == B_HI) {
            // per-limb constants of the final stage are fetched here, not held in registers all along
            const ulonglong2 cn = __ldg(reinterpret_cast<const ulonglong2*>(&Pg->ninv));
            const ulonglong2 cw = __ldg(reinterpret_cast<const ulonglong2*>(&Pg->w1ninv));
            const u64 q = q2 >> 1;
#pragma unroll
            for (int lo = 0; lo < (1 << b); ++lo) {                  // B_HI == 3 here: one group
                const int ix = lo, iy = ix + (1 << b);
                const u64 U = r[ix], V = r[iy];
                r[ix] = csub(shoup_mul(U + V, cn.x, cn.y, nq), q);
                r[iy] = csub(shoup_mul(U - V + q2, cw.x, cw.y, nq), q);
            }
        } else {
#pragma unroll
            for (int hi = 0; hi < (8 >> b); ++hi) {
                const ulonglong2 t = tw.get(b, hi);
#pragma unroll
                for (int lo = 0; lo < (1 << b); ++lo) {
                    const int ix = (hi << (b + 1)) + lo, iy = ix + (1 << b);
                    const u64 U = r[ix], V = r[iy];
                    r[ix] = csub(U + V, q2);                          // [0,2q)
                    r[iy] = shoup_mul(U - V + q2, t.x, t.y, nq);      // [0,2q)
                }
            }
        }
    }
}

// ---- shared-memory re-distribution ---------------------------------------------------------------
template <int P0>
__device__ __forceinline__ void sm_store(u64* sm, u32 tid, const u64 (&r)[kElems]) {
    if constexpr (P0 == 0) {
#pragma unroll
        for (int j = 0; j < 8; ++j)
            *reinterpret_cast<ulonglong2*>(sm + swz((tid << 4) | (2 * j))) = make_ulonglong2(r[2 * j], r[2 * j + 1]);
    } else {
#pragma unroll
        for (int i = 0; i < kElems; ++i) sm[swz(elem_index<P0>(tid, i))] = r[i];
    }
}
template <int P0>
__device__ __forceinline__ void sm_load(const u64* sm, u32 tid, u64 (&r)[kElems]) {
    if constexpr (P0 == 0) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const ulonglong2 v = *reinterpret_cast<const ulonglong2*>(sm + swz((tid << 4) | (2 * j)));
            r[2 * j] = v.x; r[2 * j + 1] = v.y;
        }
    } else {
#pragma unroll
        for (int i = 0; i < kElems; ++i) r[i] = sm[swz(elem_index<P0>(tid, i))];
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
struct TwTab {
    const ulonglong2* __restrict__ tw;   // [N]       standard bit-reversed table
    const ulonglong2* __restrict__ t0;   // [15][N/16] transposed copy for the window at bit 0
    u32 N;
};
// Twiddle source of the round whose window sits at slab bit P0; off = offset of the slab in its polynomial.
template <int P0>
__device__ __forceinline__ auto round_tw(const TwTab& T, u32 off, u32 tid) {
    if constexpr (P0 == 0) return TwT0{T.t0 + ((off >> 4) + tid), T.N >> 4};
    else return TwStd{T.tw, ((T.N + off) >> (P0 + 4)) + (tid >> P0)};
}

// ---- forward transform of a slab: the LOGS low stages --------------------------------------------
// On entry r[] holds the layout of round 0 (window at bit p0(0)); on exit the layout of the last round
// (window at bit 0: 16 contiguous residues per thread).  The slab may be a tile of several independent
// 2^LOGS-point sub-transforms (tid then spans more than 2^(LOGS-4) threads): nothing here depends on it.
template <int LOGS>
__device__ __forceinline__ void slab_fwd(u64 (&r)[kElems], u64* sm, u32 tid, const TwTab& T, u32 off, u64 q2, u64 nq) {
    using S = Sched<LOGS>;
    {
        constexpr int P0 = S::p0(0);
        fwd_round<S::top(0) - 1 - P0, 0>(r, round_tw<P0>(T, off, tid), q2, nq);
    }
    if constexpr (S::NR > 1) {
        constexpr int PP = S::p0(0), P0 = S::p0(1);
        sm_store<PP>(sm, tid, r);
        __syncthreads();
        sm_load<P0>(sm, tid, r);
        fwd_round<S::top(1) - 1 - P0, 0>(r, round_tw<P0>(T, off, tid), q2, nq);
    }
    if constexpr (S::NR > 2) {
        constexpr int PP = S::p0(1), P0 = S::p0(2);
        // no barrier needed: every thread overwrites exactly the words it read in the previous exchange
        sm_store<PP>(sm, tid, r);
        __syncthreads();
        sm_load<P0>(sm, tid, r);
        fwd_round<S::top(2) - 1 - P0, 0>(r, round_tw<P0>(T, off, tid), q2, nq);
    }
}

// ---- inverse transform of a slab: the LOGS low stages, bottom bit first --------------------------
// Entry layout: window at bit 0; exit layout: window at bit p0(0).
// LAST: the slab is the whole polynomial (N == 2^LOGS), so the top stage folds N^-1.
template <int LOGS, bool LAST>
__device__ __forceinline__ void slab_inv(u64 (&r)[kElems], u64* sm, u32 tid, const TwTab& T, u32 off, u64 q2, u64 nq,
                                         const LimbParams* __restrict__ Pg) {
    using S = Sched<LOGS>;
    if constexpr (S::NR > 2) {
        constexpr int P0 = S::p0(2), PN = S::p0(1);
        inv_round<S::top(2) - 1 - P0, 0, false>(r, round_tw<P0>(T, off, tid), q2, nq, Pg);
        sm_store<P0>(sm, tid, r);
        __syncthreads();
        sm_load<PN>(sm, tid, r);
    }
    if constexpr (S::NR > 1) {
        constexpr int P0 = S::p0(1), PN = S::p0(0);
        inv_round<S::top(1) - 1 - P0, 0, false>(r, round_tw<P0>(T, off, tid), q2, nq, Pg);
        sm_store<P0>(sm, tid, r);
        __syncthreads();
        sm_load<PN>(sm, tid, r);
    }
    {
        constexpr int P0 = S::p0(0);
        inv_round<S::top(0) - 1 - P0, 0, LAST>(r, round_tw<P0>(T, off, tid), q2, nq, Pg);
    }
}

}  // namespace gfhe
