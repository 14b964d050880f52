// sm_100a kernels for batched negacyclic NTT / INTT / fused polynomial multiplication over
// RNS limbs, layout [L, B, N] (limb outermost, N contiguous), 8-byte residues.
//
// Work unit: a TILE of 4096 residues on one 256-thread CTA (16 residues per thread in registers,
// four radix-2 stages per "round", rounds joined by an row-padded shared-memory exchange).
//
// N <= 4096 : one kernel, one CTA per limb-polynomial (the tile is the polynomial).
// N  > 4096 : two passes, N = 2^T x 2^S with S = 8 (N <= 2^16) or 12 (N <= 2^20), T = log2 N - S <= 8.
//             "column" pass = the T high stages: tile = 2^T rows (stride 2^S) x 2^(12-T) contiguous
//             columns, twiddles shared by every tile.  "row" pass = the S low stages on a contiguous
//             tile (16 independent 256-point sub-transforms when S = 8).  For N = 2^16 this is the
//             balanced 8 + 8 split: each pass does 8 stages per 16 bytes of traffic.
//             poly-mul fuses row-fwd(a), row-fwd(b), the pointwise product and row-inv into ONE
//             kernel, so no element-wise kernel ever runs; N^-1 is folded into the last inverse stage.
//
// These kernels are what arithmetic.NTT / arithmetic.iNTT / polynomial.poly_mul dispatch to;
// they replace the stubs at /root/reference/arithmetic.py:15-19 and the missing product in
// /root/reference/" polynomial.py":3-5.
// This file is a template body: ntt_harvey.cu and ntt_lazy.cu include it and instantiate Launcher<A>
// for one arithmetic policy each (modarith.cuh), so the two sets of kernels compile in parallel.
#include <atomic>
#include <cstdlib>
#include "ntt_core.cuh"
#include "pipe_sched.h"
#include "plan.h"

namespace gfhe {

// ------------------------------------------------------------------------------------------------
// row kernels: LOGS low stages on a contiguous tile of 2^LOGT residues (LOGS <= LOGT <= 12)
// ------------------------------------------------------------------------------------------------
#ifndef GFHE_ROWB
#define GFHE_ROWB 2
#endif
#ifndef GFHE_MINB32
#define GFHE_MINB32 3   // 32-bit words, where three CTAs fit in shared memory
#endif
#ifndef GFHE_MINB
#define GFHE_MINB 2   // resident 256-thread CTAs per SM the 4096-residue kernels are compiled for (128 registers)
#endif
template <int LOGT>
struct TileGeo {
    static constexpr int NT = LOGT > 4 ? (1 << (LOGT - 4)) : 1;
    static constexpr int MINB = LOGT == 12 ? GFHE_MINB : 1;   // 256-thread CTAs: cap at 128 registers (fused poly-mul kernel)
    static constexpr int ROWB = LOGT == 12 ? GFHE_ROWB : 1;   // stand-alone row passes
};

// resident CTAs per SM the stand-alone row passes are compiled for.  32-bit words need half the registers; four resident
// CTAs pay on the packed layout (2^16 NTT 0.200 -> 0.177 ms) and, once the twiddle entries are compact, on the 64-bit
// layout too (k_row_fwd 0.099 -> 0.095 ms, k_row_inv 0.095 -> 0.091 ms)
#ifndef GFHE_ROWB32
#define GFHE_ROWB32 4
#endif
template <class A, int LOGT>
struct RowOcc { static constexpr int kB = LOGT == 12 ? (sizeof(typename A::word) == 4 ? GFHE_ROWB32 : GFHE_ROWB) : 1; };

struct Tables {   // kernel argument: the plan's device tables
    const ulonglong2* tw;    // [L][N]
    const ulonglong2* t0;    // [L][N/512][15][32]
    const LimbParams* params;
};

struct TileId {
    u64 g0;     // flat offset of the tile in the [L,B,N] array
    u32 off;    // offset of the tile inside its polynomial
    u32 limb;
};
// ---- L2 prefetch of a later CTA's input ------------------------------------------------------------------------------
// CTAs are dispatched in blockIdx order and two per SM are resident, so the CTA that takes over this one's slot is about
// blockIdx.x + 296.  Asking L2 for its input tile now (one instruction, no registers) turns its cold HBM reads at start-up,
// when it has nothing else to do, into L2 hits.  Measured at config-4 shape (profiles/r01_run11_prefetch.txt):
//   k_row_polymul  60-bit 0.550 -> 0.531 ms, 50-bit 0.358 -> 0.343 ms   (distance 148 slightly better than 296)
//   k_row_fwd      30-bit 0.165 -> 0.142 ms, 50-bit -1 %, 60-bit +2 % (so not used by the 60-bit policies)
//   column kernels +2..15 % slower with a per-line prefetch.global.L2: not used there.
#ifndef GFHE_PF_MUL
#define GFHE_PF_MUL 148
#endif
#ifndef GFHE_PF_ROW
#define GFHE_PF_ROW 296
#endif
__device__ __forceinline__ void l2_prefetch_bulk(const void* p, u32 bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p), "r"(bytes));
}
template <int LOGT, u32 DIST, class S>
__device__ __forceinline__ void prefetch_row_tile(const S* base, u32 tid) {
    if (DIST != 0 && LOGT >= 4 && tid == 0 && blockIdx.x + DIST < gridDim.x)
        l2_prefetch_bulk(base + ((u64)(blockIdx.x + DIST) << LOGT), (u32)sizeof(S) << LOGT);
}
template <class A> struct RowPf {   // k_row_inv: 30-bit inverse step 0.221 -> 0.207 ms, 50-bit 0.250 -> 0.269 ms (so 32-bit words only)
    static constexpr u32 kFwd = (sizeof(typename A::word) == 4 || A::kFloat) ? GFHE_PF_ROW : 0;
    static constexpr u32 kInv = sizeof(typename A::word) == 4 ? GFHE_PF_ROW : 0;
};

template <int LOGT>
__device__ __forceinline__ TileId tile_id(u32 tile, u32 logN, u32 B, u32 limb0) {
    TileId s;
    s.g0 = (u64)tile << LOGT;
    const u32 p = (u32)(s.g0 >> logN);
    s.off = (u32)s.g0 & ((1u << logN) - 1u);
    s.limb = limb0 + p / B;
    return s;
}
template <class A>
__device__ __forceinline__ TwTab<typename A::tw_t> limb_tables(const Tables& T, u32 limb, u32 logN) {
    typedef typename A::tw_t E;
    const size_t N = (size_t)1 << logN;
    return TwTab<E>{reinterpret_cast<const E*>(T.tw) + limb * N, reinterpret_cast<const E*>(T.t0) + limb * t0_entries(N), (u32)N};
}

// Per-limb scalars every kernel needs: Cf/Ci = offset constants of the arithmetic policy.
template <class W> struct LimbK { W q, nq, Cf, Ci; };
template <class A>
__device__ __forceinline__ LimbK<typename A::word> limb_consts(const LimbParams* __restrict__ Pg) {
    typedef typename A::word W;
    if constexpr (A::kFloat) {   // FP64 policy: Cf = Ci = q, nq = 1/q (ntt_core.cuh)
        const double q = __ldg(&Pg->qd), qi = __ldg(&Pg->qinv);
        return LimbK<W>{q, qi, q, q};
    } else {
    const ulonglong2 v = __ldg(reinterpret_cast<const ulonglong2*>(&Pg->q2));   // {2q, -q mod 2^64}
    return LimbK<W>{(W)(v.x >> 1), (W)v.y, (W)((u64)(A::kFwdC / 2) * v.x), (W)((u64)(A::kInvC / 2) * v.x)};
    }
}
template <class A>
__device__ __forceinline__ void fwd_final_all(typename A::word (&r)[kElems], const LimbK<typename A::word>& K, const LimbParams* __restrict__ Pg) {
    u32 ks = 0, km = 0;
    if constexpr (A::kLazy) { ks = __ldg(&Pg->kshift); km = (u32)__ldg(&Pg->kmu); }
#pragma unroll
    for (int i = 0; i < kElems; ++i) r[i] = fwd_final<A>(r[i], K.q, K.nq, ks, km);
}

template <class A>
__device__ __forceinline__ void fwd_semi_all(typename A::word (&r)[kElems], const LimbK<typename A::word>& K, const LimbParams* __restrict__ Pg) {
    u32 ks = 0, km = 0;
    if constexpr (A::kLazy) { ks = __ldg(&Pg->kshift); km = (u32)__ldg(&Pg->kmu); }
#pragma unroll
    for (int i = 0; i < kElems; ++i) r[i] = fwd_semi<A>(r[i], K.q, K.nq, ks, km);
}

// CG: the words were written by another CTA of the same launch (pipelined kernels below): read them from L2, never from
// this SM's L1, which may still hold what the same addresses held earlier in the launch.
template <bool CG, class T>
__device__ __forceinline__ T ld_res(const T* p) { if constexpr (CG) return __ldcg(p); else return *p; }

template <int LOGT>
__device__ __forceinline__ void copy_out_padded(const u64* sm, u64* out, u32 tid) {
    constexpr int NT = TileGeo<LOGT>::NT;
#pragma unroll
    for (int it = 0; it < 8; ++it) {
        const u32 e = 2u * (it * NT + tid);
        *reinterpret_cast<ulonglong2*>(out + e) = *reinterpret_cast<const ulonglong2*>(sm + pad(e));
    }
}
// FP64 policy: canonical doubles in shared memory <-> canonical 64-bit integers in HBM
template <int LOGT>
__device__ __forceinline__ void copy_out_padded(const double* sm, u64* out, u32 tid) {
    constexpr int NT = TileGeo<LOGT>::NT;
#pragma unroll
    for (int it = 0; it < 8; ++it) {
        const u32 e = 2u * (it * NT + tid);
        const double2 v = *reinterpret_cast<const double2*>(sm + pad(e));
        *reinterpret_cast<ulonglong2*>(out + e) = make_ulonglong2(Fp50::to_u64(v.x), Fp50::to_u64(v.y));
    }
}
template <int LOGT>
__device__ __forceinline__ void copy_in_padded(double* sm, const u64* in, u32 tid) {
    constexpr int NT = TileGeo<LOGT>::NT;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        ulonglong2 v[4];
#pragma unroll
        for (int it = 0; it < 4; ++it) v[it] = *reinterpret_cast<const ulonglong2*>(in + 2u * ((4 * h + it) * NT + tid));
#pragma unroll
        for (int it = 0; it < 4; ++it)
            *reinterpret_cast<double2*>(sm + pad(2u * ((4 * h + it) * NT + tid))) = make_double2(Fp50::from_u64(v[it].x), Fp50::from_u64(v[it].y));
    }
}
// 32-bit words in shared memory, 64-bit words in HBM: pad(e) is even for even e, so the pair is one 8-byte access
template <int LOGT>
__device__ __forceinline__ void copy_out_padded(const u32* sm, u64* out, u32 tid) {
    constexpr int NT = TileGeo<LOGT>::NT;
#pragma unroll
    for (int it = 0; it < 8; ++it) {
        const u32 e = 2u * (it * NT + tid);
        const uint2 v = *reinterpret_cast<const uint2*>(sm + pad(e));
        *reinterpret_cast<ulonglong2*>(out + e) = make_ulonglong2(v.x, v.y);
    }
}
template <int LOGT>
__device__ __forceinline__ void copy_in_padded(u32* sm, const u64* in, u32 tid) {
    constexpr int NT = TileGeo<LOGT>::NT;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        ulonglong2 v[4];
#pragma unroll
        for (int it = 0; it < 4; ++it) v[it] = *reinterpret_cast<const ulonglong2*>(in + 2u * ((4 * h + it) * NT + tid));
#pragma unroll
        for (int it = 0; it < 4; ++it) *reinterpret_cast<uint2*>(sm + pad(2u * ((4 * h + it) * NT + tid))) = make_uint2((u32)v[it].x, (u32)v[it].y);
    }
}
// 32-bit words in shared memory and in HBM: four residues (one 16-byte global access) never straddle a padded row
template <int LOGT>
__device__ __forceinline__ void copy_out_padded(const u32* sm, u32* out, u32 tid) {
    constexpr int NT = TileGeo<LOGT>::NT;
#pragma unroll
    for (int it = 0; it < 4; ++it) {
        const u32 e = 4u * (it * NT + tid);
        const uint2 lo = *reinterpret_cast<const uint2*>(sm + pad(e)), hi = *reinterpret_cast<const uint2*>(sm + pad(e) + 2);
        *reinterpret_cast<uint4*>(out + e) = make_uint4(lo.x, lo.y, hi.x, hi.y);
    }
}
template <int LOGT>
__device__ __forceinline__ void copy_in_padded(u32* sm, const u32* in, u32 tid) {
    constexpr int NT = TileGeo<LOGT>::NT;
    uint4 v[4];
#pragma unroll
    for (int it = 0; it < 4; ++it) v[it] = *reinterpret_cast<const uint4*>(in + 4u * (it * NT + tid));
#pragma unroll
    for (int it = 0; it < 4; ++it) {
        u32* d = sm + pad(4u * (it * NT + tid));
        *reinterpret_cast<uint2*>(d) = make_uint2(v[it].x, v[it].y);
        *reinterpret_cast<uint2*>(d + 2) = make_uint2(v[it].z, v[it].w);
    }
}
template <int LOGT>
__device__ __forceinline__ void copy_in_padded(u64* sm, const u64* in, u32 tid) {
    constexpr int NT = TileGeo<LOGT>::NT;
#pragma unroll
    for (int h = 0; h < 2; ++h) {   // two batches of four 16-byte loads in flight
        ulonglong2 v[4];
#pragma unroll
        for (int it = 0; it < 4; ++it) v[it] = *reinterpret_cast<const ulonglong2*>(in + 2u * ((4 * h + it) * NT + tid));
#pragma unroll
        for (int it = 0; it < 4; ++it) *reinterpret_cast<ulonglong2*>(sm + pad(2u * ((4 * h + it) * NT + tid))) = v[it];
    }
}

// ---- TMA bulk copy (cp.async.bulk) of one contiguous tile into shared memory, completion on an mbarrier ----
__device__ __forceinline__ u32 smem_u32(const void* p) { return (u32)__cvta_generic_to_shared(p); }
// one thread: arm the barrier with the byte count and start the copy
// Shared-memory layout of the fused poly-mul kernels, in bytes: exchange buffer | stash of NTT(a) (the persistent form
// also lands the next a tile there, linear, in storage words) | b's landing buffer | two mbarriers.
template <class A, int LOGT, bool PERSIST>
struct MulSmem {
    typedef typename A::word W;
    typedef typename A::store S;
    static constexpr u32 kExch = ((u32)sizeof(W) * padded_words(LOGT) + 15u) & ~15u;
    static constexpr u32 kTile = (u32)sizeof(S) << LOGT;
    static constexpr u32 kStash = (PERSIST && kTile > kExch) ? kTile : kExch;
    static constexpr u32 kOff1 = kExch, kOffB = kExch + kStash, kOffBar = kOffB + kTile, kBytes = kOffBar + 16;
    static_assert(kOff1 % 16 == 0 && kOffB % 16 == 0 && kOffBar % 8 == 0, "TMA destinations are 16-byte aligned");
    // resident CTAs per SM the kernel is compiled for: three when shared memory (227 KB) and 80 registers allow
    static constexpr int kFit = (int)((227u * 1024u) / (kBytes + 1024u));
    static constexpr int kMinB = LOGT == 12 ? ((sizeof(W) == 4 && kFit >= 3) ? (kFit < GFHE_MINB32 ? kFit : GFHE_MINB32) : GFHE_MINB) : 1;
};

__device__ __forceinline__ void tma_tile_load(void* dst_smem, const void* src_gmem, u32 bytes, u64* mbar) {
    const u32 bar = smem_u32(mbar), dst = smem_u32(dst_smem);
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src_gmem), "r"(bytes), "r"(bar) : "memory");
}
// every thread: wait for phase 0 of the barrier (the copy has landed and is visible)
__device__ __forceinline__ void tma_tile_wait(u64* mbar) {
    const u32 bar = smem_u32(mbar);
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "GFHE_TMA_WAIT:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n\t"
        "@!p bra GFHE_TMA_WAIT;\n\t"
        "}\n" ::"r"(bar) : "memory");
}

// Forward: the LOGS low stages + full reduction to [0, q).  Always the last pass of a forward NTT.
// The tile bodies are device functions of the tile index: the stand-alone kernels call them with blockIdx.x, the pipelined
// kernels at the end of this file with the tile their ticket names.
template <class A, int LOGS, int LOGT, bool CG = false>
__device__ __forceinline__ void row_fwd_body(const typename A::store* in, typename A::store* out, const Tables& tabs, u32 logN, u32 B, u32 limb0,
                                             u32 tile, typename A::word* sm) {
    typedef typename A::word W;
    typedef typename A::store S;
    const u32 tid = threadIdx.x;
    const TileId s = tile_id<LOGT>(tile, logN, B, limb0);
    const LimbParams* __restrict__ Pg = tabs.params + s.limb;
    const LimbK<W> K = limb_consts<A>(Pg);
    const auto T = limb_tables<A>(tabs, s.limb, logN);
    W r[kElems];
    constexpr int P0 = Sched<LOGS>::p0(0);
#ifdef GFHE_EXP_TMAROW
    // experiment ("TMA staging of every sub-transform"): the tile arrives by ONE cp.async.bulk into the
    // (not yet used) exchange buffer instead of 16 LDG per thread, and the registers are filled from shared memory
    if constexpr (LOGT == 12 && sizeof(S) == sizeof(W) && !A::kFloat) {
        __shared__ __align__(8) u64 exp_mbar;
        if (tid == 0) tma_tile_load(sm, in + s.g0, (u32)sizeof(S) << LOGT, &exp_mbar);
        __syncthreads();
        tma_tile_wait(&exp_mbar);
#pragma unroll
        for (int i = 0; i < kElems; ++i) r[i] = A::from_raw(reinterpret_cast<const S*>(sm)[elem_index<P0>(tid, 0) + (i << P0)]);
        __syncthreads();   // the exchange of the first round overwrites the landing buffer
    } else
#endif
    if (A::kFloat && logN > (u32)LOGT) {   // second pass of a large transform: the column pass left raw words (they differ from canonical ones for the FP64 policy only)
#pragma unroll
        for (int i = 0; i < kElems; ++i) r[i] = GFHE_GLD(A::from_raw, ld_res<CG>(in + s.g0 + elem_index<P0>(tid, 0) + (i << P0)), tid * 77u + i);
    } else {
#pragma unroll
        for (int i = 0; i < kElems; ++i) r[i] = GFHE_GLD(A::from_u64, ld_res<CG>(in + s.g0 + elem_index<P0>(tid, 0) + (i << P0)), tid * 77u + i);
    }
    slab_fwd<A, LOGS>(r, sm, tid, T, s.off, K.Cf, K.nq);
    fwd_final_all<A>(r, K, Pg);
#ifdef GFHE_EXP_DIRECTST   // experiment: 16-byte stores straight from the registers (each lane writes its own 128-byte row)
    if constexpr (LOGT > 4) {
        S* po = out + s.g0 + (tid << 4);
#pragma unroll
        for (int j = 0; j < 8; ++j) *reinterpret_cast<ulonglong2*>(po + 2 * j) = make_ulonglong2(A::to_u64(r[2 * j]), A::to_u64(r[2 * j + 1]));
    } else {
#else
    if constexpr (LOGT > 4) {
        sm_store<0>(sm, tid, r);   // own 128-byte row: no hazard with the preceding sm_load<0>
        GFHE_SYNC();
        copy_out_padded<LOGT>(sm, out + s.g0, tid);
    } else {
#endif
#pragma unroll
        for (int i = 0; i < kElems; ++i) out[s.g0 + i] = A::to_u64(r[i]);
    }
}

template <class A, int LOGS, int LOGT>
__global__ void __launch_bounds__(TileGeo<LOGT>::NT, RowOcc<A, LOGT>::kB)
k_row_fwd(const typename A::store* in, typename A::store* out, const Tables tabs, u32 logN, u32 B, u32 limb0) {
    typedef typename A::word W;
    __shared__ __align__(16) W sm[padded_words(LOGT)];
    prefetch_row_tile<LOGT, RowPf<A>::kFwd>(in, threadIdx.x);
    row_fwd_body<A, LOGS, LOGT>(in, out, tabs, logN, B, limb0, blockIdx.x, sm);
}

// Inverse: the LOGS low stages (bottom first).  Always the first pass of an inverse NTT; when the
// tile is the whole polynomial it also folds N^-1 and fully reduces.
template <class A, int LOGS, int LOGT, bool WHOLE>
__device__ __forceinline__ void row_inv_body(const typename A::store* in, typename A::store* out, const Tables& tabs, u32 logN, u32 B, u32 limb0,
                                             u32 tile, typename A::word* sm) {
    typedef typename A::word W;
    const u32 tid = threadIdx.x;
    const TileId s = tile_id<LOGT>(tile, logN, B, limb0);
    const LimbParams* __restrict__ Pg = tabs.params + s.limb;
    const LimbK<W> K = limb_consts<A>(Pg);
    const auto T = limb_tables<A>(tabs, s.limb, logN);
    W r[kElems];
    if constexpr (LOGT > 4) {
        copy_in_padded<LOGT>(sm, in + s.g0, tid);
        GFHE_SYNC();
        sm_load<0>(sm, tid, r);
    } else {
#pragma unroll
        for (int i = 0; i < kElems; ++i) r[i] = A::from_u64(in[s.g0 + i]);
    }
    slab_inv<A, LOGS, WHOLE, true>(r, sm, tid, T, s.off, K.Ci, K.nq, Pg);   // inputs of an inverse transform are canonical
    constexpr int P0 = Sched<LOGS>::p0(0);
#pragma unroll
    for (int i = 0; i < kElems; ++i) (out + s.g0 + elem_index<P0>(tid, 0))[i << P0] = WHOLE ? A::to_u64(r[i]) : A::to_raw(r[i]);
}

template <class A, int LOGS, int LOGT, bool WHOLE>
__global__ void __launch_bounds__(TileGeo<LOGT>::NT, RowOcc<A, LOGT>::kB)
k_row_inv(const typename A::store* in, typename A::store* out, const Tables tabs, u32 logN, u32 B, u32 limb0) {
    typedef typename A::word W;
    __shared__ __align__(16) W sm[padded_words(LOGT)];
    prefetch_row_tile<LOGT, RowPf<A>::kInv>(in, threadIdx.x);
    row_inv_body<A, LOGS, LOGT, WHOLE>(in, out, tabs, logN, B, limb0, blockIdx.x, sm);
}

// mbarrier pieces used by the pipelined and persistent kernels: init once, then one phase per tile
__device__ __forceinline__ void mbar_init(u64* mbar) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(mbar)));
}
__device__ __forceinline__ void tma_tile_issue(void* dst_smem, const void* src_gmem, u32 bytes, u64* mbar) {
    const u32 bar = smem_u32(mbar), dst = smem_u32(dst_smem);
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src_gmem), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(u64* mbar, u32 parity) {
    const u32 bar = smem_u32(mbar);
    u32 done;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}\n"
                     : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    } while (!done);
}

// Fused product on a tile: c = row-inv( row-fwd(a) .* row-fwd(b) ).  For N <= 4096 (WHOLE) this
// is the complete negacyclic product of one limb-polynomial in a single launch.
// BNTT: b is already in the NTT domain (canonical residues, the order gfhe_ntt_forward writes): its forward transform is
// skipped and NTT(a) never leaves the registers -- a third of the butterflies gone for callers that cache NTT(b).
// PIPE: called from a pipelined kernel (end of this file): a and b were written by other CTAs of the same launch (a is read
// through L2 only; the TMA copy of b is ordered after the dependency wait by the caller), the mbarrier was initialised once
// at kernel start and this tile uses its phase *parity (a shared-memory word, read where the wait is).
template <class A, int LOGS, int LOGT, bool WHOLE, bool BNTT, bool PIPE>
__device__ __forceinline__ void row_polymul_body(const typename A::store* a, const typename A::store* b, typename A::store* c, const Tables& tf,
                                                 const Tables& ti, u32 logN, u32 B, u32 limb0, u32 tile, unsigned char* smem, const volatile u32* parity) {
    typedef typename A::word W;
    typedef typename A::store S;
    using Lay = MulSmem<A, LOGT, false>;
    W* sm0 = reinterpret_cast<W*>(smem);                           // exchange buffer
    W* sm1 = reinterpret_cast<W*>(smem + Lay::kOff1);              // stash for NTT(a) while NTT(b) is computed
    S* smb = reinterpret_cast<S*>(smem + Lay::kOffB);              // b's tile, fetched by one TMA bulk copy while a is transformed
    u64* mbar = reinterpret_cast<u64*>(smem + Lay::kOffBar);
    const u32 tid = threadIdx.x;
    const TileId s = tile_id<LOGT>(tile, logN, B, limb0);
    if constexpr (PIPE) {
        if (tid == 0) tma_tile_issue(smb, b + s.g0, (u32)sizeof(S) << LOGT, mbar);
    } else {
        if (tid == 0) tma_tile_load(smb, b + s.g0, (u32)sizeof(S) << LOGT, mbar);
        prefetch_row_tile<LOGT, GFHE_PF_MUL>(a, tid);
        prefetch_row_tile<LOGT, GFHE_PF_MUL>(b, tid);
    }
    const LimbParams* __restrict__ Pg = tf.params + s.limb;
    const LimbK<W> K = limb_consts<A>(Pg);
    const auto Tf = limb_tables<A>(tf, s.limb, logN);
    const auto Ti = limb_tables<A>(ti, s.limb, logN);
    constexpr int P0 = Sched<LOGS>::p0(0);
    W r[kElems];
#pragma unroll
    for (int i = 0; i < kElems; ++i)
        r[i] = WHOLE ? GFHE_GLD(A::from_u64, ld_res<PIPE>(a + s.g0 + elem_index<P0>(tid, 0) + (i << P0)), tid * 77u + i)
                     : GFHE_GLD(A::from_raw, ld_res<PIPE>(a + s.g0 + elem_index<P0>(tid, 0) + (i << P0)), tid * 77u + i);
    slab_fwd<A, LOGS, WHOLE ? 1 : 16>(r, sm0, tid, Tf, s.off, K.Cf, K.nq);   // whole polynomial: canonical inputs
    fwd_semi_all<A>(r, K, Pg);
    if constexpr (BNTT) {
        if constexpr (PIPE) mbar_wait(mbar, *parity); else tma_tile_wait(mbar);
        const S* pb = smb + (tid << 4);   // this thread's 16 residues of NTT(b), same positions as r[]
        if constexpr (A::kFloat) {
#pragma unroll
            for (int i = 0; i < kElems; ++i) {
                const double vb = fp_reduce(A::from_u64(pb[i]), K.q, K.nq);   // canonical [0,q) -> within 0.51q
                r[i] = fp_mulmod(r[i], vb, vb * K.nq, K.q);
            }
        } else if constexpr (sizeof(W) == 8) {
            const u64 mu = __ldg(&Pg->bar_mu);
            const u32 bs = __ldg(&Pg->bar_shift);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const ulonglong2 vb = *reinterpret_cast<const ulonglong2*>(pb + 2 * j);
                if constexpr (A::kLazy) {
                    r[2 * j] = barrett_lazy(r[2 * j], vb.x, K.nq, mu, bs);
                    r[2 * j + 1] = barrett_lazy(r[2 * j + 1], vb.y, K.nq, mu, bs);
                } else {
                    r[2 * j] = barrett_mul(r[2 * j], vb.x, K.q, mu, bs);
                    r[2 * j + 1] = barrett_mul(r[2 * j + 1], vb.y, K.q, mu, bs);
                }
            }
        } else {
            const u64 b64 = __ldg(&Pg->bar64);
#pragma unroll
            for (int i = 0; i < kElems; ++i) {
                const u64 z = (u64)r[i] * (W)pb[i];
                const u64 qh = mulhi64(z, b64);
                r[i] = csub((W)z + (W)qh * K.nq, K.q);
            }
        }
        GFHE_SYNC();   // the exchange buffer is reused by the inverse transform
    } else {
        sm_store<0>(sm1, tid, r);   // private row of sm1: written and read back by this thread only
        GFHE_SYNC();            // every thread has finished reading a's last exchange out of sm0 (and sees the barrier)
        if constexpr (PIPE) mbar_wait(mbar, *parity); else tma_tile_wait(mbar);
    #pragma unroll
        for (int i = 0; i < kElems; ++i) r[i] = WHOLE ? A::from_u64((smb + elem_index<P0>(tid, 0))[i << P0]) : A::from_raw((smb + elem_index<P0>(tid, 0))[i << P0]);
        slab_fwd<A, LOGS, WHOLE ? 1 : 16>(r, sm0, tid, Tf, s.off, K.Cf, K.nq);   // whole polynomial: canonical inputs
        fwd_semi_all<A>(r, K, Pg);
        if constexpr (A::kFloat) {   // |a|, |b| <= 0.51q: the same error-free product, quotient by 1/q
            const W* pa = sm1 + pad(tid << 4);
#pragma unroll
            for (int i = 0; i < kElems; ++i) r[i] = fp_mulmod(pa[i], r[i], r[i] * K.nq, K.q);
        } else if constexpr (sizeof(W) == 8) {
            const u64 mu = __ldg(&Pg->bar_mu);
            const u32 bs = __ldg(&Pg->bar_shift);
            const u64* pa = reinterpret_cast<const u64*>(sm1) + pad(tid << 4);
    #pragma unroll
            for (int j = 0; j < 8; ++j) {
                const ulonglong2 va = *reinterpret_cast<const ulonglong2*>(pa + 2 * j);
                if constexpr (A::kLazy) {
                    r[2 * j] = barrett_lazy(va.x, r[2 * j], K.nq, mu, bs);
                    r[2 * j + 1] = barrett_lazy(va.y, r[2 * j + 1], K.nq, mu, bs);
                } else {
                    r[2 * j] = barrett_mul(va.x, r[2 * j], K.q, mu, bs);
                    r[2 * j + 1] = barrett_mul(va.y, r[2 * j + 1], K.q, mu, bs);
                }
            }
        } else {   // 32-bit words: a, b < q < 2^30, one-word Barrett on the 60-bit product
            const u64 b64 = __ldg(&Pg->bar64);
            const W* pa = sm1 + pad(tid << 4);
    #pragma unroll
            for (int i = 0; i < kElems; ++i) {
                const u64 z = (u64)pa[i] * r[i];
                const u64 qh = mulhi64(z, b64);                  // floor(z/q) - 1 <= qh <= floor(z/q)
                r[i] = csub((W)z + (W)qh * K.nq, K.q);           // z - qh*q in [0, 2q) fits 32 bits
            }
        }
    }
    slab_inv<A, LOGS, WHOLE>(r, sm0, tid, Ti, s.off, K.Ci, K.nq, Pg);
#pragma unroll
    for (int i = 0; i < kElems; ++i) (c + s.g0 + elem_index<P0>(tid, 0))[i << P0] = WHOLE ? A::to_u64(r[i]) : A::to_raw(r[i]);
}

template <class A, int LOGS, int LOGT, bool WHOLE, bool BNTT = false>
__global__ void __launch_bounds__(TileGeo<LOGT>::NT, MulSmem<A, LOGT, false>::kMinB)
k_row_polymul(const typename A::store* a, const typename A::store* b, typename A::store* c, const Tables tf, const Tables ti, u32 logN, u32 B,
              u32 limb0) {
    extern __shared__ __align__(16) unsigned char smem[];
    row_polymul_body<A, LOGS, LOGT, WHOLE, BNTT, false>(a, b, c, tf, ti, logN, B, limb0, blockIdx.x, smem, nullptr);
}

// ---- persistent form of the fused row pass (two-pass transforms only) ---------------------------------------------------
// One CTA per resident slot loops over tiles; the NEXT tile's operands are fetched by TMA bulk copies while the current
// tile is computed, so no CTA ever waits for HBM at its start:
//   a' -> the stash buffer (free once the pointwise product has read NTT(a)), landing during the inverse transform
//   b' -> b's landing buffer (free once b has been read), landing during fwd(b), the product and the inverse transform
// mbarrier phase `it & 1` of barA / barB belongs to loop trip `it`.
template <class A, int LOGS, int LOGT>
__global__ void __launch_bounds__(TileGeo<LOGT>::NT, MulSmem<A, LOGT, true>::kMinB)
k_row_polymul_persist(const typename A::store* a, const typename A::store* b, typename A::store* c, const Tables tf, const Tables ti, u32 logN,
                      u32 B, u32 limb0, u32 ntiles) {
    typedef typename A::word W;
    typedef typename A::store S;
    using Lay = MulSmem<A, LOGT, true>;
    extern __shared__ __align__(16) unsigned char smem[];
    W* sm0 = reinterpret_cast<W*>(smem);                           // exchange buffer
    S* sa = reinterpret_cast<S*>(smem + Lay::kOff1);               // a's landing buffer (linear), then the stash of NTT(a)
    W* sm1 = reinterpret_cast<W*>(sa);
    S* smb = reinterpret_cast<S*>(smem + Lay::kOffB);              // b's landing buffer
    u64* barA = reinterpret_cast<u64*>(smem + Lay::kOffBar);
    u64* barB = barA + 1;
    const u32 tid = threadIdx.x;
    constexpr u32 kTileBytes = (u32)sizeof(S) << LOGT;
    constexpr int P0 = Sched<LOGS>::p0(0);
    if (tid == 0) {
        mbar_init(barA); mbar_init(barB);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        const u64 g0 = (u64)blockIdx.x << LOGT;
        tma_tile_issue(sa, a + g0, kTileBytes, barA);
        tma_tile_issue(smb, b + g0, kTileBytes, barB);
    }
    __syncthreads();
    u32 it = 0;
#pragma unroll 1
    for (u32 tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
        const u32 par = it & 1u;
        const u32 next = tile + gridDim.x;
        TileId s;
        s.g0 = (u64)tile << LOGT;
        s.off = (u32)s.g0 & ((1u << logN) - 1u);
        s.limb = limb0 + (u32)(s.g0 >> logN) / B;
        const LimbParams* __restrict__ Pg = tf.params + s.limb;
        const LimbK<W> K = limb_consts<A>(Pg);
        const auto Tf = limb_tables<A>(tf, s.limb, logN);
        W r[kElems];
        mbar_wait(barA, par);
#pragma unroll
        for (int i = 0; i < kElems; ++i) r[i] = A::from_raw((sa + elem_index<P0>(tid, 0))[i << P0]);
        slab_fwd<A, LOGS>(r, sm0, tid, Tf, s.off, K.Cf, K.nq);   // its exchange barrier also orders every read of sa before the stash
        fwd_semi_all<A>(r, K, Pg);
        sm_store<0>(sm1, tid, r);
        __syncthreads();
        mbar_wait(barB, par);
#pragma unroll
        for (int i = 0; i < kElems; ++i) r[i] = A::from_raw((smb + elem_index<P0>(tid, 0))[i << P0]);
        // per-phase constants are re-derived from an opaque copy of the limb index: kept live from the top of the trip (or,
        // for twiddles, from a's transform into b's) they spill ~40 registers under the 128-register cap
        u32 limb_b = s.limb;
        asm volatile("" : "+r"(limb_b));
        slab_fwd<A, LOGS>(r, sm0, tid, limb_tables<A>(tf, limb_b, logN), s.off, K.Cf, K.nq);   // ... and every read of smb before the refill below
        if (tid == 0 && next < ntiles) tma_tile_issue(smb, b + ((u64)next << LOGT), kTileBytes, barB);
        fwd_semi_all<A>(r, K, Pg);
        if constexpr (A::kFloat) {
            const W* pa = sm1 + pad(tid << 4);
#pragma unroll
            for (int i = 0; i < kElems; ++i) r[i] = fp_mulmod(pa[i], r[i], r[i] * K.nq, K.q);
        } else if constexpr (sizeof(W) == 8) {
            const u64 mu = __ldg(&Pg->bar_mu);
            const u32 bs = __ldg(&Pg->bar_shift);
            const u64* pa = reinterpret_cast<const u64*>(sm1) + pad(tid << 4);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const ulonglong2 va = *reinterpret_cast<const ulonglong2*>(pa + 2 * j);
                if constexpr (A::kLazy) {
                    r[2 * j] = barrett_lazy(va.x, r[2 * j], K.nq, mu, bs);
                    r[2 * j + 1] = barrett_lazy(va.y, r[2 * j + 1], K.nq, mu, bs);
                } else {
                    r[2 * j] = barrett_mul(va.x, r[2 * j], K.q, mu, bs);
                    r[2 * j + 1] = barrett_mul(va.y, r[2 * j + 1], K.q, mu, bs);
                }
            }
        } else {
            const u64 b64 = __ldg(&Pg->bar64);
            const W* pa = sm1 + pad(tid << 4);
#pragma unroll
            for (int i = 0; i < kElems; ++i) {
                const u64 z = (u64)pa[i] * r[i];
                const u64 qh = mulhi64(z, b64);
                r[i] = csub((W)z + (W)qh * K.nq, K.q);
            }
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // the stash was written by generic stores; TMA writes next
        __syncthreads();                                                // every thread is done with the stash and with sm0
        if (tid == 0 && next < ntiles) tma_tile_issue(sa, a + ((u64)next << LOGT), kTileBytes, barA);
        u32 limb_i = s.limb;
        asm volatile("" : "+r"(limb_i));
        const LimbParams* __restrict__ Pi = tf.params + limb_i;
        const LimbK<W> Ki = limb_consts<A>(Pi);
        slab_inv<A, LOGS, false>(r, sm0, tid, limb_tables<A>(ti, limb_i, logN), s.off, Ki.Ci, Ki.nq, Pi);
#pragma unroll
        for (int i = 0; i < kElems; ++i) (c + s.g0 + elem_index<P0>(tid, 0))[i << P0] = A::to_raw(r[i]);
        __syncthreads();   // sm0 is reused by the next trip's forward exchange
    }
}

// ------------------------------------------------------------------------------------------------
// column kernels (N > 4096): tile = 2^T rows x C columns, T = 12 - LOGC, C = 2^LOGC, row stride 2^S;
// tile index e = k*C + j  <->  coefficient k*2^S + tile*C + j.  Transform bits of e: [LOGC, 12).
// Twiddles depend only on the row bits, so every tile of a limb reads the same 2^T - 1 entries.
// S is a template parameter (8 or 12) so every register's global offset is an immediate.
// ------------------------------------------------------------------------------------------------
template <int LOGC, int S>
struct ColGeo {
    static constexpr int LO8 = LOGC > 8 ? LOGC - 8 : 0;   // lowest local bit of the window at bit 8
    static constexpr int LO4 = LOGC > 4 ? LOGC - 4 : 0;   // lowest local bit of the window at bit 4
    static constexpr bool TWO = LOGC < 8;                  // needs the second window
    static constexpr int LOGN = S + 12 - LOGC, LT = S - LOGC;   // 2^LT tiles per polynomial
    // offset of tile element e from the tile base
    __device__ static __forceinline__ u32 off(u32 e) { return (e & ((1u << LOGC) - 1u)) + ((e >> LOGC) << S); }
    // off() moves bit fields, and the thread bits and register bits of elem_index<P0>(tid, i) are disjoint, so
    // off(elem_index<P0>(tid, i)) = toff<P0>(tid) + roff<P0>(i) with roff a compile-time constant (an immediate).
    __host__ __device__ static constexpr u32 off_c(u32 e) { return (e & ((1u << LOGC) - 1u)) + ((e >> LOGC) << S); }
    template <int P0> __device__ static __forceinline__ u32 toff(u32 tid) { return off(elem_index<P0>(tid, 0)); }
    template <int P0> __host__ __device__ static constexpr u32 roff(int i) { return off_c((u32)i << P0); }
};
struct ColId { u64 base; u32 limb; };
template <class W> constexpr int kColPad = sizeof(W) == 4 ? 1 : 2;   // padding of the column kernels' exchange (ntt_core.cuh, padk)
template <int LOGC, int S>
__device__ __forceinline__ ColId col_id(u32 flat, u32 B, u32 limb0) {   // flat = polynomial * 2^LT + tile
    using Geo = ColGeo<LOGC, S>;
    const u32 p = flat >> Geo::LT, tile = flat & ((1u << Geo::LT) - 1u);
    return ColId{((u64)p << Geo::LOGN) + ((u64)tile << LOGC), limb0 + p / B};
}

#ifndef GFHE_COLB
#define GFHE_COLB 4   // 64 registers: four resident CTAs per SM (measured: k_col_fwd -2.7 %, k_col_inv -8.7 % against 2 -> 3 resident)
#endif
// Exchange buffer (only when the pass has two windows) and the staged twiddles of a column tile.
template <class A, int LOGC> struct ColSmem {
    static constexpr u32 kExchWords = (LOGC < 8) ? padded_words(12) : 2;
};

template <class A, int LOGC, int S>
__device__ __forceinline__ void col_fwd_body(const typename A::store* in0, typename A::store* out0, const Tables& tabs, u32 B, u32 limb0, u32 flat,
                                             typename A::word* sm, typename A::tw_t* tws) {
    using Geo = ColGeo<LOGC, S>;
    typedef typename A::word W;
    typedef typename A::store St;
    typedef typename A::tw_t E;
    const u32 tid = threadIdx.x;
    const ColId id = col_id<LOGC, S>(flat, B, limb0);
    const St* in = in0 + id.base;
    St* out = out0 + id.base;
    const LimbParams* __restrict__ Pg = tabs.params + id.limb;
    const LimbK<W> K = limb_consts<A>(Pg);
    W r[kElems];
    {
        const St* p = in + Geo::template toff<8>(tid);
#pragma unroll
        for (int i = 0; i < kElems; ++i) r[i] = GFHE_GLD(A::from_u64, p[Geo::template roff<8>(i)], tid * 77u + i);
    }
    tws[tid] = __ldg(reinterpret_cast<const E*>(tabs.tw) + ((size_t)id.limb << Geo::LOGN) + tid);   // the column pass reads entries [1, 2^T) only
    GFHE_SYNC();
    fwd_round<A, 3, Geo::LO8, TwSm<E>, 1>(r, TwSm<E>{tws, 1u}, K.Cf, K.nq);   // canonical inputs: the range schedule starts from q
    if constexpr (Geo::TWO) {
        sm_store<8, W, kColPad<W>>(sm, tid, r);
        GFHE_SYNC();
        sm_load<4, W, kColPad<W>>(sm, tid, r);
        fwd_round<A, 3, Geo::LO4, TwSm<E>, fwd_bound(1, 4 - Geo::LO8)>(r, TwSm<E>{tws, 16u + (tid >> 4)}, K.Cf, K.nq);
        St* p = out + Geo::template toff<4>(tid);
#pragma unroll
        for (int i = 0; i < kElems; ++i) p[Geo::template roff<4>(i)] = A::to_raw(r[i]);
    } else {
        St* p = out + Geo::template toff<8>(tid);
#pragma unroll
        for (int i = 0; i < kElems; ++i) p[Geo::template roff<8>(i)] = A::to_raw(r[i]);
    }
}

template <class A, int LOGC, int S>
__global__ void __launch_bounds__(256, GFHE_COLB)
k_col_fwd(const typename A::store* in0, typename A::store* out0, const typename A::store* in1, typename A::store* out1, const Tables tabs,
          u32 B, u32 limb0) {
    __shared__ __align__(16) typename A::word sm[ColSmem<A, LOGC>::kExchWords];
    __shared__ typename A::tw_t tws[256];
    col_fwd_body<A, LOGC, S>(blockIdx.y ? in1 : in0, blockIdx.y ? out1 : out0, tabs, B, limb0, blockIdx.x, sm, tws);
}

// CG: the input is the output of row tiles of the same launch (pipelined kernels)
template <class A, int LOGC, int S, bool CG = false>
__device__ __forceinline__ void col_inv_body(const typename A::store* in0, typename A::store* out0, const Tables& tabs, u32 B, u32 limb0, u32 flat,
                                             typename A::word* sm, typename A::tw_t* tws) {
    using Geo = ColGeo<LOGC, S>;
    typedef typename A::word W;
    typedef typename A::store St;
    typedef typename A::tw_t E;
    const u32 tid = threadIdx.x;
    const ColId id = col_id<LOGC, S>(flat, B, limb0);
    const St* in = in0 + id.base;
    St* out = out0 + id.base;
    const LimbParams* __restrict__ Pg = tabs.params + id.limb;
    const LimbK<W> K = limb_consts<A>(Pg);
    W r[kElems];
    if constexpr (Geo::TWO) {
        const St* p = in + Geo::template toff<4>(tid);
#pragma unroll
        for (int i = 0; i < kElems; ++i) r[i] = GFHE_GLD(A::from_raw, ld_res<CG>(p + Geo::template roff<4>(i)), tid * 77u + i);
    } else {
        const St* p = in + Geo::template toff<8>(tid);
#pragma unroll
        for (int i = 0; i < kElems; ++i) r[i] = GFHE_GLD(A::from_raw, ld_res<CG>(p + Geo::template roff<8>(i)), tid * 77u + i);
    }
    tws[tid] = __ldg(reinterpret_cast<const E*>(tabs.tw) + ((size_t)id.limb << Geo::LOGN) + tid);
    GFHE_SYNC();
    if constexpr (Geo::TWO) {
        inv_round<A, 3, Geo::LO4, false>(r, TwSm<E>{tws, 16u + (tid >> 4)}, K.Ci, K.nq, Pg);
        sm_store<4, W, kColPad<W>>(sm, tid, r);
        GFHE_SYNC();
        sm_load<8, W, kColPad<W>>(sm, tid, r);
    }
    inv_round<A, 3, Geo::LO8, true>(r, TwSm<E>{tws, 1u}, K.Ci, K.nq, Pg);
    St* p = out + Geo::template toff<8>(tid);
#pragma unroll
    for (int i = 0; i < kElems; ++i) p[Geo::template roff<8>(i)] = A::to_u64(r[i]);
}

template <class A, int LOGC, int S>
__global__ void __launch_bounds__(256, GFHE_COLB)
k_col_inv(const typename A::store* in0, typename A::store* out0, const Tables tabs, u32 B, u32 limb0) {
    __shared__ __align__(16) typename A::word sm[ColSmem<A, LOGC>::kExchWords];
    __shared__ typename A::tw_t tws[256];
    col_inv_body<A, LOGC, S>(in0, out0, tabs, B, limb0, blockIdx.x, sm, tws);
}

// ------------------------------------------------------------------------------------------------
// host-side launchers, one set per arithmetic policy
// ------------------------------------------------------------------------------------------------
#define GFHE_ROW_CASES(X) X(4) X(5) X(6) X(7) X(8) X(9) X(10) X(11) X(12)
#define GFHE_COL_CASES(X) X(4) X(5) X(6) X(7)

static inline Tables fwd_tables(const Plan& pl) { return Tables{pl.d_fwd, pl.d_fwd_t0, pl.d_params}; }
static inline Tables inv_tables(const Plan& pl) { return Tables{pl.d_inv, pl.d_inv_t0, pl.d_params}; }

// Stage split of a two-pass transform: S low (row) stages, log2 N - S <= 8 high (column) stages.
static inline u32 row_stages(u32 logN) { return logN <= 16 ? 8u : 12u; }

#include "ntt_pipe.inl"   // inside namespace gfhe

// N > 4096: the pipelined single launch when the plan asks for it (GFHE_PIPE=1 at plan creation); default: one launch per pass
#ifndef GFHE_PIPE_DEFAULT
#define GFHE_PIPE_DEFAULT 0
#endif
static inline bool use_pipe(const Plan& pl) { return pl.pipe < 0 ? GFHE_PIPE_DEFAULT != 0 : pl.pipe == 1; }

static inline unsigned tiles_of(const Plan& pl, u32 nl, u32 B) {
    const u32 logT = pl.logN < (u32)kMaxLogSlab ? pl.logN : (u32)kMaxLogSlab;
    return (unsigned)(((u64)nl * B) << (pl.logN - logT));
}

template <class A, int LOGS, int LOGT, bool WHOLE, bool BNTT = false>
static cudaError_t launch_row_polymul(const Plan& pl, u32 limb0, const typename A::store* a, const typename A::store* b, typename A::store* c, u32 B, unsigned nt,
                                      cudaStream_t st) {
    const size_t smem = MulSmem<A, LOGT, false>::kBytes;
    static std::atomic<bool> configured[64];   // per template instantiation, per device (zero-initialised; setting the attribute twice is harmless)
    const int dev = pl.device & 63;
    if (smem > 48 * 1024 && !configured[dev].load(std::memory_order_acquire)) {
        cudaError_t e = cudaFuncSetAttribute(k_row_polymul<A, LOGS, LOGT, WHOLE, BNTT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        configured[dev].store(true, std::memory_order_release);
    }
    { ProfScope ps(pl, K_ROW_POLYMUL, st); k_row_polymul<A, LOGS, LOGT, WHOLE, BNTT><<<nt, TileGeo<LOGT>::NT, smem, st>>>(a, b, c, fwd_tables(pl), inv_tables(pl), pl.logN, B, limb0); }
    return cudaGetLastError();
}

// The persistent form wins where it compiles without spills (32-bit words: k_row_polymul 0.316 -> 0.278 ms at config-4
// shape); the 64-bit-word policies spill ~40 registers under the 128-register cap and lose 9-13 %.  GFHE_PERSIST=0/1 (read when the plan is created) overrides.
template <class A>
static inline bool use_persistent(const Plan& pl) { return pl.persist < 0 ? sizeof(typename A::word) == 4 : pl.persist == 1; }
template <class A, int LOGS, int LOGT>
static cudaError_t launch_row_polymul_persist(const Plan& pl, u32 limb0, const typename A::store* a, const typename A::store* b, typename A::store* c, u32 B, unsigned nt,
                                              cudaStream_t st) {
    const size_t smem = MulSmem<A, LOGT, true>::kBytes;
    static std::atomic<int> slots[64];   // resident CTAs per device, 0 = not configured yet (per template instantiation)
    const int dev = pl.device & 63;
    int nslots = slots[dev].load(std::memory_order_acquire);
    if (nslots == 0) {
        cudaError_t e = cudaFuncSetAttribute(k_row_polymul_persist<A, LOGS, LOGT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        int sms = 148, per_sm = 1;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, pl.device);
        if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_row_polymul_persist<A, LOGS, LOGT>, TileGeo<LOGT>::NT, smem)) != cudaSuccess) return e;
        nslots = sms * (per_sm > 0 ? per_sm : 1);
        slots[dev].store(nslots, std::memory_order_release);
    }
    const unsigned grid = nt < (unsigned)nslots ? nt : (unsigned)nslots;
    { ProfScope ps(pl, K_ROW_POLYMUL, st);
      k_row_polymul_persist<A, LOGS, LOGT><<<grid, TileGeo<LOGT>::NT, smem, st>>>(a, b, c, fwd_tables(pl), inv_tables(pl), pl.logN, B, limb0, nt); }
    return cudaGetLastError();
}

// column pass of a forward transform: in0 -> out0 and (if in1) in1 -> out1 in one launch
template <class A>
static cudaError_t launch_col_fwd(const Plan& pl, u32 limb0, const typename A::store* in0, typename A::store* out0, const typename A::store* in1,
                                  typename A::store* out1, u32 B,
                                  unsigned nt, cudaStream_t st) {
    const u32 S = row_stages(pl.logN);
    const dim3 grid(nt, in1 ? 2 : 1);
    ProfScope ps(pl, K_COL_FWD, st);
    switch ((12 - (int)(pl.logN - S)) * 16 + (int)S) {
#define X(LC) case LC * 16 + 8: k_col_fwd<A, LC, 8><<<grid, 256, 0, st>>>(in0, out0, in1, out1, fwd_tables(pl), B, limb0); break; \
              case LC * 16 + 12: k_col_fwd<A, LC, 12><<<grid, 256, 0, st>>>(in0, out0, in1, out1, fwd_tables(pl), B, limb0); break;
        GFHE_COL_CASES(X)
#undef X
        default: return cudaErrorInvalidValue;
    }
    return cudaSuccess;
}
template <class A>
static cudaError_t launch_col_inv(const Plan& pl, u32 limb0, const typename A::store* in, typename A::store* out, u32 B, unsigned nt, cudaStream_t st) {
    const u32 S = row_stages(pl.logN);
    ProfScope ps(pl, K_COL_INV, st);
    switch ((12 - (int)(pl.logN - S)) * 16 + (int)S) {
#define X(LC) case LC * 16 + 8: k_col_inv<A, LC, 8><<<nt, 256, 0, st>>>(in, out, inv_tables(pl), B, limb0); break; \
              case LC * 16 + 12: k_col_inv<A, LC, 12><<<nt, 256, 0, st>>>(in, out, inv_tables(pl), B, limb0); break;
        GFHE_COL_CASES(X)
#undef X
        default: return cudaErrorInvalidValue;
    }
    return cudaSuccess;
}

template <class A>
cudaError_t Launcher<A>::ntt_fwd(const Plan& pl, u32 limb0, u32 nl, const typename A::store* in, typename A::store* out, u32 B, cudaStream_t st) {
    const unsigned nt = tiles_of(pl, nl, B);
    if (pl.logN <= (u32)kMaxLogSlab) {
        switch (pl.logN) {
#define X(LS) case LS: { ProfScope ps(pl, K_ROW_FWD, st); k_row_fwd<A, LS, LS><<<nt, TileGeo<LS>::NT, 0, st>>>(in, out, fwd_tables(pl), pl.logN, B, limb0); } break;
            GFHE_ROW_CASES(X)
#undef X
            default: return cudaErrorInvalidValue;
        }
        return cudaGetLastError();
    }
    if (use_pipe(pl)) return pipe_ntt<A>(pl, false, limb0, nl, in, out, B, st);
    if (cudaError_t e = launch_col_fwd<A>(pl, limb0, in, out, nullptr, nullptr, B, nt, st)) return e;
    ProfScope ps(pl, K_ROW_FWD, st);
    if (row_stages(pl.logN) == 8) k_row_fwd<A, 8, 12><<<nt, 256, 0, st>>>(out, out, fwd_tables(pl), pl.logN, B, limb0);
    else k_row_fwd<A, 12, 12><<<nt, 256, 0, st>>>(out, out, fwd_tables(pl), pl.logN, B, limb0);
    return cudaGetLastError();
}

template <class A>
cudaError_t Launcher<A>::ntt_inv(const Plan& pl, u32 limb0, u32 nl, const typename A::store* in, typename A::store* out, u32 B, cudaStream_t st) {
    const unsigned nt = tiles_of(pl, nl, B);
    if (pl.logN <= (u32)kMaxLogSlab) {
        switch (pl.logN) {
#define X(LS) case LS: { ProfScope ps(pl, K_ROW_INV, st); k_row_inv<A, LS, LS, true><<<nt, TileGeo<LS>::NT, 0, st>>>(in, out, inv_tables(pl), pl.logN, B, limb0); } break;
            GFHE_ROW_CASES(X)
#undef X
            default: return cudaErrorInvalidValue;
        }
        return cudaGetLastError();
    }
    if (use_pipe(pl)) return pipe_ntt<A>(pl, true, limb0, nl, in, out, B, st);
    {
        ProfScope ps(pl, K_ROW_INV, st);
        if (row_stages(pl.logN) == 8) k_row_inv<A, 8, 12, false><<<nt, 256, 0, st>>>(in, out, inv_tables(pl), pl.logN, B, limb0);
        else k_row_inv<A, 12, 12, false><<<nt, 256, 0, st>>>(in, out, inv_tables(pl), pl.logN, B, limb0);
    }
    if (cudaError_t e = launch_col_inv<A>(pl, limb0, out, out, B, nt, st)) return e;
    return cudaGetLastError();
}

template <class A>
cudaError_t Launcher<A>::poly_mul(const Plan& pl, u32 limb0, u32 nl, const typename A::store* a, const typename A::store* b, typename A::store* c, typename A::store* ws, u32 B,
                                  cudaStream_t st) {
    const unsigned nt = tiles_of(pl, nl, B);
    if (pl.logN <= (u32)kMaxLogSlab) {
        switch (pl.logN) {
#define X(LS) case LS: return launch_row_polymul<A, LS, LS, true>(pl, limb0, a, b, c, B, nt, st);
            GFHE_ROW_CASES(X)
#undef X
            default: return cudaErrorInvalidValue;
        }
    }
    if (!ws) return cudaErrorInvalidValue;
    if (use_pipe(pl)) return pipe_polymul<A, false>(pl, limb0, nl, a, b, c, ws, B, st);
    // high stages of a -> c (used as scratch), of b -> ws, in one launch (grid.y picks the operand)
    if (cudaError_t e = launch_col_fwd<A>(pl, limb0, a, c, b, ws, B, nt, st)) return e;
    cudaError_t e;
    if (row_stages(pl.logN) == 8 && use_persistent<A>(pl)) e = launch_row_polymul_persist<A, 8, 12>(pl, limb0, c, ws, c, B, nt, st);
    else e = row_stages(pl.logN) == 8 ? launch_row_polymul<A, 8, 12, false>(pl, limb0, c, ws, c, B, nt, st)
                                      : launch_row_polymul<A, 12, 12, false>(pl, limb0, c, ws, c, B, nt, st);
    if (e != cudaSuccess) return e;
    if ((e = launch_col_inv<A>(pl, limb0, c, c, B, nt, st)) != cudaSuccess) return e;
    return cudaGetLastError();
}

template <class A>
cudaError_t Launcher<A>::poly_mul_pre(const Plan& pl, u32 limb0, u32 nl, const typename A::store* a, const typename A::store* b_ntt, typename A::store* c, u32 B,
                                      cudaStream_t st) {
    const unsigned nt = tiles_of(pl, nl, B);
    if (pl.logN <= (u32)kMaxLogSlab) {
        switch (pl.logN) {
#define X(LS) case LS: return launch_row_polymul<A, LS, LS, true, true>(pl, limb0, a, b_ntt, c, B, nt, st);
            GFHE_ROW_CASES(X)
#undef X
            default: return cudaErrorInvalidValue;
        }
    }
    if (use_pipe(pl)) return pipe_polymul<A, true>(pl, limb0, nl, a, b_ntt, c, nullptr, B, st);
    // high stages of a -> c (scratch), then the fused row pass against NTT(b), then the high inverse stages in place
    if (cudaError_t e = launch_col_fwd<A>(pl, limb0, a, c, nullptr, nullptr, B, nt, st)) return e;
    cudaError_t e = row_stages(pl.logN) == 8 ? launch_row_polymul<A, 8, 12, false, true>(pl, limb0, c, b_ntt, c, B, nt, st)
                                             : launch_row_polymul<A, 12, 12, false, true>(pl, limb0, c, b_ntt, c, B, nt, st);
    if (e != cudaSuccess) return e;
    if ((e = launch_col_inv<A>(pl, limb0, c, c, B, nt, st)) != cudaSuccess) return e;
    return cudaGetLastError();
}

}  // namespace gfhe
