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

__device__ __forceinline__ ulonglong2 ldg_tw(const ulonglong2* p) { return __ldg(p); }

// ---- one round of forward (CT) stages on local bits B_HI..B_LO --------------------------------
// v = (twiddle-table index of the thread's group at the round's top stage); the table index of
// local bit b, sub-group hi is (v << (3-b)) + hi  (see DESIGN.md "twiddle indexing").
template <int B_HI, int B_LO>
__device__ __forceinline__ void fwd_round(u64 (&r)[kElems], const ulonglong2* __restrict__ tw, u32 v,
                                          u64 q2, u64 nq) {
#pragma unroll
    for (int b = B_HI; b >= B_LO; --b) {
#pragma unroll
        for (int hi = 0; hi < (8 >> b); ++hi) {
            const ulonglong2 t = ldg_tw(tw + ((v << (3 - b)) + hi));
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
template <int B_HI, int B_LO, bool LAST>
__device__ __forceinline__ void inv_round(u64 (&r)[kElems], const ulonglong2* __restrict__ tw, u32 v,
                                          const LimbParams& P) {
    const u64 q2 = P.q2, nq = P.nq;
#pragma unroll
    for (int b = B_LO; b <= B_HI; ++b) {
        if (LAST && b == B_HI) {
#pragma unroll
            for (int lo = 0; lo < (1 << b); ++lo) {                  // B_HI == 3 here: one group
                const int ix = lo, iy = ix + (1 << b);
                const u64 U = r[ix], V = r[iy];
                r[ix] = csub(shoup_mul(U + V, P.ninv, P.ninv_s, nq), P.q);
                r[iy] = csub(shoup_mul(U - V + q2, P.w1ninv, P.w1ninv_s, nq), P.q);
            }
        } else {
#pragma unroll
            for (int hi = 0; hi < (8 >> b); ++hi) {
                const ulonglong2 t = ldg_tw(tw + ((v << (3 - b)) + hi));
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

// ---- forward transform of a contiguous slab: all LOGS low stages ----------------------------------
// On entry r[] holds the layout of round 0 (window at bit p0(0)); on exit the layout of the
// last round (window at bit 0: 16 contiguous residues per thread).  G = N + (offset of the slab
// inside its polynomial).
template <int LOGS>
__device__ __forceinline__ void slab_fwd(u64 (&r)[kElems], u64* sm, u32 tid, const ulonglong2* __restrict__ tw,
                                         u32 G, u64 q2, u64 nq) {
    using S = Sched<LOGS>;
    {
        constexpr int P0 = S::p0(0);
        fwd_round<S::top(0) - 1 - P0, 0>(r, tw, (G >> (P0 + 4)) + (tid >> P0), q2, nq);
    }
    if constexpr (S::NR > 1) {
        constexpr int PP = S::p0(0), P0 = S::p0(1);
        sm_store<PP>(sm, tid, r);
        __syncthreads();
        sm_load<P0>(sm, tid, r);
        fwd_round<S::top(1) - 1 - P0, 0>(r, tw, (G >> (P0 + 4)) + (tid >> P0), q2, nq);
    }
    if constexpr (S::NR > 2) {
        constexpr int PP = S::p0(1), P0 = S::p0(2);
        sm_store<PP>(sm, tid, r);
        __syncthreads();
        sm_load<P0>(sm, tid, r);
        fwd_round<S::top(2) - 1 - P0, 0>(r, tw, (G >> (P0 + 4)) + (tid >> P0), q2, nq);
    }
}

// ---- inverse transform of a contiguous slab: the LOGS low stages, bottom bit first ----------------
// Entry layout: window at bit p0(NR-1) = 0; exit layout: window at bit p0(0).
// LAST: the slab is the whole polynomial (N == 2^LOGS), so the top stage folds N^-1.
template <int LOGS, bool LAST>
__device__ __forceinline__ void slab_inv(u64 (&r)[kElems], u64* sm, u32 tid, const ulonglong2* __restrict__ tw,
                                         u32 G, const LimbParams& P) {
    using S = Sched<LOGS>;
    if constexpr (S::NR > 2) {
        constexpr int P0 = S::p0(2), PN = S::p0(1);
        inv_round<S::top(2) - 1 - P0, 0, false>(r, tw, (G >> (P0 + 4)) + (tid >> P0), P);
        sm_store<P0>(sm, tid, r);
        __syncthreads();
        sm_load<PN>(sm, tid, r);
    }
    if constexpr (S::NR > 1) {
        constexpr int P0 = S::p0(1), PN = S::p0(0);
        inv_round<S::top(1) - 1 - P0, 0, false>(r, tw, (G >> (P0 + 4)) + (tid >> P0), P);
        sm_store<P0>(sm, tid, r);
        __syncthreads();
        sm_load<PN>(sm, tid, r);
    }
    {
        constexpr int P0 = S::p0(0);
        inv_round<S::top(0) - 1 - P0, 0, LAST>(r, tw, (G >> (P0 + 4)) + (tid >> P0), P);
    }
}

}  // namespace gfhe
