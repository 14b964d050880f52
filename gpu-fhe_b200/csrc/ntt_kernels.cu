// sm_100a kernels for batched negacyclic NTT / INTT / fused polynomial multiplication over
// RNS limbs, layout [L, B, N] (limb outermost, N contiguous), 8-byte residues.
//
// N <= 4096 : one kernel, one CTA per limb-polynomial (the slab is the polynomial).
// N  > 4096 : N = N1 x 4096.  "column" kernel = the log2(N1) high stages on stride-4096
//             columns (twiddles shared by every column), "row" kernel = the 12 low stages on
//             contiguous 4096-residue rows.  poly-mul fuses  row-fwd(a), row-fwd(b), pointwise
//             product and row-inv into ONE kernel, so no element-wise kernel ever runs.
//
// These kernels are what arithmetic.NTT / arithmetic.iNTT / polynomial.poly_mul dispatch to;
// they replace the stubs at /root/reference/arithmetic.py:15-19 and the missing product in
// /root/reference/" polynomial.py":3-5.
#include "ntt_core.cuh"
#include "plan.h"

namespace gfhe {

// ------------------------------------------------------------------------------------------------
// row kernels: contiguous slab of 2^LOGS residues, LOGS = min(log2 N, 12)
// ------------------------------------------------------------------------------------------------
template <int LOGS>
struct RowGeo {
    static constexpr int NT = LOGS > 4 ? (1 << (LOGS - 4)) : 1;
};

struct SlabId {
    u64 g0;     // flat offset of the slab in the [L,B,N] array
    u32 off;    // offset of the slab inside its polynomial
    u32 limb;
};
template <int LOGS>
__device__ __forceinline__ SlabId slab_id(u32 logN, u32 B, u32 limb0) {
    SlabId s;
    s.g0 = (u64)blockIdx.x << LOGS;
    const u32 p = (u32)(s.g0 >> logN);
    s.off = (u32)s.g0 & ((1u << logN) - 1u);
    s.limb = limb0 + p / B;
    return s;
}

template <int LOGS>
__device__ __forceinline__ void copy_out_swizzled(const u64* sm, u64* out, u32 tid) {
    constexpr int NT = RowGeo<LOGS>::NT;
#pragma unroll
    for (int it = 0; it < 8; ++it) {
        const u32 e = 2u * (it * NT + tid);
        *reinterpret_cast<ulonglong2*>(out + e) = *reinterpret_cast<const ulonglong2*>(sm + swz(e));
    }
}
template <int LOGS>
__device__ __forceinline__ void copy_in_swizzled(u64* sm, const u64* in, u32 tid) {
    constexpr int NT = RowGeo<LOGS>::NT;
    ulonglong2 v[8];
#pragma unroll
    for (int it = 0; it < 8; ++it) v[it] = *reinterpret_cast<const ulonglong2*>(in + 2u * (it * NT + tid));
#pragma unroll
    for (int it = 0; it < 8; ++it) *reinterpret_cast<ulonglong2*>(sm + swz(2u * (it * NT + tid))) = v[it];
}

// Forward: the LOGS low stages + full reduction to [0, q).  Always the last pass of a forward NTT.
template <int LOGS>
__global__ void __launch_bounds__(RowGeo<LOGS>::NT)
k_row_fwd(const u64* in, u64* out, const ulonglong2* __restrict__ twf,
          const LimbParams* __restrict__ params, u32 logN, u32 B, u32 limb0) {
    __shared__ __align__(16) u64 sm[1 << LOGS];
    const u32 tid = threadIdx.x;
    const SlabId s = slab_id<LOGS>(logN, B, limb0);
    const LimbParams P = params[s.limb];
    const ulonglong2* tw = twf + ((size_t)s.limb << logN);
    u64 r[kElems];
    constexpr int P0 = Sched<LOGS>::p0(0);
#pragma unroll
    for (int i = 0; i < kElems; ++i) r[i] = in[s.g0 + elem_index<P0>(tid, i)];
    slab_fwd<LOGS>(r, sm, tid, tw, (1u << logN) + s.off, P.q2, P.nq);
#pragma unroll
    for (int i = 0; i < kElems; ++i) r[i] = csub(csub(r[i], P.q2), P.q);
    if constexpr (LOGS > 4) {
        sm_store<0>(sm, tid, r);   // own 128-byte row: no hazard with the preceding sm_load<0>
        __syncthreads();
        copy_out_swizzled<LOGS>(sm, out + s.g0, tid);
    } else {
#pragma unroll
        for (int i = 0; i < kElems; ++i) out[s.g0 + i] = r[i];
    }
}

// Inverse: the LOGS low stages (bottom first).  Always the first pass of an inverse NTT; when the
// slab is the whole polynomial it also folds N^-1 and fully reduces.
template <int LOGS, bool WHOLE>
__global__ void __launch_bounds__(RowGeo<LOGS>::NT)
k_row_inv(const u64* in, u64* out, const ulonglong2* __restrict__ twi,
          const LimbParams* __restrict__ params, u32 logN, u32 B, u32 limb0) {
    __shared__ __align__(16) u64 sm[1 << LOGS];
    const u32 tid = threadIdx.x;
    const SlabId s = slab_id<LOGS>(logN, B, limb0);
    const LimbParams P = params[s.limb];
    const ulonglong2* tw = twi + ((size_t)s.limb << logN);
    u64 r[kElems];
    if constexpr (LOGS > 4) {
        copy_in_swizzled<LOGS>(sm, in + s.g0, tid);
        __syncthreads();
        sm_load<0>(sm, tid, r);
    } else {
#pragma unroll
        for (int i = 0; i < kElems; ++i) r[i] = in[s.g0 + i];
    }
    slab_inv<LOGS, WHOLE>(r, sm, tid, tw, (1u << logN) + s.off, P);
    constexpr int P0 = Sched<LOGS>::p0(0);
#pragma unroll
    for (int i = 0; i < kElems; ++i) out[s.g0 + elem_index<P0>(tid, i)] = r[i];
}

// Fused product on a slab: c = row-inv( row-fwd(a) .* row-fwd(b) ).  For N <= 4096 (WHOLE) this
// is the complete negacyclic product of one limb-polynomial in a single launch.
template <int LOGS, bool WHOLE>
__global__ void __launch_bounds__(RowGeo<LOGS>::NT)
k_row_polymul(const u64* a, const u64* b, u64* c,
              const ulonglong2* __restrict__ twf, const ulonglong2* __restrict__ twi,
              const LimbParams* __restrict__ params, u32 logN, u32 B, u32 limb0) {
    extern __shared__ __align__(16) u64 smem[];
    u64* sm0 = smem;                 // exchange buffer
    u64* sm1 = smem + (1 << LOGS);   // stash for NTT(a) while NTT(b) is computed
    const u32 tid = threadIdx.x;
    const SlabId s = slab_id<LOGS>(logN, B, limb0);
    const LimbParams P = params[s.limb];
    const ulonglong2* tf = twf + ((size_t)s.limb << logN);
    const ulonglong2* ti = twi + ((size_t)s.limb << logN);
    const u32 G = (1u << logN) + s.off;
    constexpr int P0 = Sched<LOGS>::p0(0);
    u64 r[kElems];
#pragma unroll
    for (int i = 0; i < kElems; ++i) r[i] = a[s.g0 + elem_index<P0>(tid, i)];
    slab_fwd<LOGS>(r, sm0, tid, tf, G, P.q2, P.nq);
#pragma unroll
    for (int i = 0; i < kElems; ++i) r[i] = csub(csub(r[i], P.q2), P.q);
    sm_store<0>(sm1, tid, r);
    __syncthreads();   // every thread has finished reading a's last exchange out of sm0
#pragma unroll
    for (int i = 0; i < kElems; ++i) r[i] = b[s.g0 + elem_index<P0>(tid, i)];
    slab_fwd<LOGS>(r, sm0, tid, tf, G, P.q2, P.nq);
    {
        u64 ra[kElems];
        sm_load<0>(sm1, tid, ra);
#pragma unroll
        for (int i = 0; i < kElems; ++i) r[i] = barrett_mul(ra[i], csub(csub(r[i], P.q2), P.q), P);
    }
    slab_inv<LOGS, WHOLE>(r, sm0, tid, ti, G, P);
#pragma unroll
    for (int i = 0; i < kElems; ++i) c[s.g0 + elem_index<P0>(tid, i)] = r[i];
}

// ------------------------------------------------------------------------------------------------
// column kernels (N > 4096): slab = N1 rows x C columns, N1 = 2^(12-LOGC), C = 2^LOGC; slab index
// e = k*C + j  <->  coefficient k*4096 + tile*C + j.  Transform bits of e: [LOGC, 12).
// ------------------------------------------------------------------------------------------------
template <int LOGC>
struct ColGeo {
    static constexpr int LO8 = LOGC > 8 ? LOGC - 8 : 0;   // lowest local bit of the window at bit 8
    static constexpr int LO4 = LOGC > 4 ? LOGC - 4 : 0;   // lowest local bit of the window at bit 4
    static constexpr bool TWO = LOGC < 8;                  // needs the second window
    __device__ static __forceinline__ u64 addr(u64 base, u32 e) {
        return base + (e & ((1u << LOGC) - 1u)) + ((u64)(e >> LOGC) << kMaxLogSlab);
    }
};

template <int LOGC>
__global__ void __launch_bounds__(256)
k_col_fwd(const u64* in0, u64* out0, const u64* in1, u64* out1,
          const ulonglong2* __restrict__ twf, const LimbParams* __restrict__ params, u32 B, u32 limb0) {
    using Geo = ColGeo<LOGC>;
    constexpr u32 logN = 24 - LOGC;
    __shared__ __align__(16) u64 sm[Geo::TWO ? 4096 : 2];
    const u64* in = blockIdx.y ? in1 : in0;
    u64* out = blockIdx.y ? out1 : out0;
    const u32 tid = threadIdx.x;
    const u32 p = blockIdx.x >> (12 - LOGC), tile = blockIdx.x & ((1u << (12 - LOGC)) - 1u);
    const u64 base = ((u64)p << logN) + ((u64)tile << LOGC);
    const u32 limb = limb0 + p / B;
    const LimbParams P = params[limb];
    const ulonglong2* tw = twf + ((size_t)limb << logN);
    u64 r[kElems];
#pragma unroll
    for (int i = 0; i < kElems; ++i) r[i] = in[Geo::addr(base, elem_index<8>(tid, i))];
    fwd_round<3, Geo::LO8>(r, tw, 1u, P.q2, P.nq);
    if constexpr (Geo::TWO) {
        sm_store<8>(sm, tid, r);
        __syncthreads();
        sm_load<4>(sm, tid, r);
        fwd_round<3, Geo::LO4>(r, tw, 16u + (tid >> 4), P.q2, P.nq);
#pragma unroll
        for (int i = 0; i < kElems; ++i) out[Geo::addr(base, elem_index<4>(tid, i))] = r[i];
    } else {
#pragma unroll
        for (int i = 0; i < kElems; ++i) out[Geo::addr(base, elem_index<8>(tid, i))] = r[i];
    }
}

template <int LOGC>
__global__ void __launch_bounds__(256)
k_col_inv(const u64* in, u64* out, const ulonglong2* __restrict__ twi,
          const LimbParams* __restrict__ params, u32 B, u32 limb0) {
    using Geo = ColGeo<LOGC>;
    constexpr u32 logN = 24 - LOGC;
    __shared__ __align__(16) u64 sm[Geo::TWO ? 4096 : 2];
    const u32 tid = threadIdx.x;
    const u32 p = blockIdx.x >> (12 - LOGC), tile = blockIdx.x & ((1u << (12 - LOGC)) - 1u);
    const u64 base = ((u64)p << logN) + ((u64)tile << LOGC);
    const u32 limb = limb0 + p / B;
    const LimbParams P = params[limb];
    const ulonglong2* tw = twi + ((size_t)limb << logN);
    u64 r[kElems];
    if constexpr (Geo::TWO) {
#pragma unroll
        for (int i = 0; i < kElems; ++i) r[i] = in[Geo::addr(base, elem_index<4>(tid, i))];
        inv_round<3, Geo::LO4, false>(r, tw, 16u + (tid >> 4), P);
        sm_store<4>(sm, tid, r);
        __syncthreads();
        sm_load<8>(sm, tid, r);
    } else {
#pragma unroll
        for (int i = 0; i < kElems; ++i) r[i] = in[Geo::addr(base, elem_index<8>(tid, i))];
    }
    inv_round<3, Geo::LO8, true>(r, tw, 1u, P);
#pragma unroll
    for (int i = 0; i < kElems; ++i) out[Geo::addr(base, elem_index<8>(tid, i))] = r[i];
}

// ------------------------------------------------------------------------------------------------
// host-side dispatch
// ------------------------------------------------------------------------------------------------
#define GFHE_ROW_CASES(X) X(4) X(5) X(6) X(7) X(8) X(9) X(10) X(11) X(12)
#define GFHE_COL_CASES(X) X(4) X(5) X(6) X(7) X(8) X(9) X(10) X(11)

static inline unsigned slabs_of(const Plan& pl, u32 nl, u32 B) {
    const u32 logS = pl.logN < (u32)kMaxLogSlab ? pl.logN : (u32)kMaxLogSlab;
    return (unsigned)(((u64)nl * B) << (pl.logN - logS));
}

cudaError_t launch_ntt_fwd(const Plan& pl, u32 limb0, u32 nl, const u64* in, u64* out, u32 B, cudaStream_t st) {
    const unsigned nslab = slabs_of(pl, nl, B);
    if (pl.logN <= (u32)kMaxLogSlab) {
        switch (pl.logN) {
#define X(LS) case LS: { ProfScope ps(pl, K_ROW_FWD, st); k_row_fwd<LS><<<nslab, RowGeo<LS>::NT, 0, st>>>(in, out, pl.d_fwd, pl.d_params, pl.logN, B, limb0); } break;
            GFHE_ROW_CASES(X)
#undef X
            default: return cudaErrorInvalidValue;
        }
    } else {
        switch (24 - (int)pl.logN) {
#define X(LC) case LC: { ProfScope ps(pl, K_COL_FWD, st); k_col_fwd<LC><<<dim3(nslab, 1), 256, 0, st>>>(in, out, nullptr, nullptr, pl.d_fwd, pl.d_params, B, limb0); } break;
            GFHE_COL_CASES(X)
#undef X
            default: return cudaErrorInvalidValue;
        }
        { ProfScope ps(pl, K_ROW_FWD, st); k_row_fwd<12><<<nslab, 256, 0, st>>>(out, out, pl.d_fwd, pl.d_params, pl.logN, B, limb0); }
    }
    return cudaGetLastError();
}

cudaError_t launch_ntt_inv(const Plan& pl, u32 limb0, u32 nl, const u64* in, u64* out, u32 B, cudaStream_t st) {
    const unsigned nslab = slabs_of(pl, nl, B);
    if (pl.logN <= (u32)kMaxLogSlab) {
        switch (pl.logN) {
#define X(LS) case LS: { ProfScope ps(pl, K_ROW_INV, st); k_row_inv<LS, true><<<nslab, RowGeo<LS>::NT, 0, st>>>(in, out, pl.d_inv, pl.d_params, pl.logN, B, limb0); } break;
            GFHE_ROW_CASES(X)
#undef X
            default: return cudaErrorInvalidValue;
        }
    } else {
        { ProfScope ps(pl, K_ROW_INV, st); k_row_inv<12, false><<<nslab, 256, 0, st>>>(in, out, pl.d_inv, pl.d_params, pl.logN, B, limb0); }
        switch (24 - (int)pl.logN) {
#define X(LC) case LC: { ProfScope ps(pl, K_COL_INV, st); k_col_inv<LC><<<nslab, 256, 0, st>>>(out, out, pl.d_inv, pl.d_params, B, limb0); } break;
            GFHE_COL_CASES(X)
#undef X
            default: return cudaErrorInvalidValue;
        }
    }
    return cudaGetLastError();
}

template <int LOGS, bool WHOLE>
static cudaError_t launch_row_polymul(const Plan& pl, u32 limb0, const u64* a, const u64* b, u64* c, u32 B, unsigned nslab,
                                      cudaStream_t st) {
    const size_t smem = (size_t)2 * sizeof(u64) << LOGS;
    static bool configured[64] = {};   // per template instantiation, per device
    const int dev = pl.device & 63;
    if (!configured[dev] && smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(k_row_polymul<LOGS, WHOLE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        configured[dev] = true;
    }
    { ProfScope ps(pl, K_ROW_POLYMUL, st); k_row_polymul<LOGS, WHOLE><<<nslab, RowGeo<LOGS>::NT, smem, st>>>(a, b, c, pl.d_fwd, pl.d_inv, pl.d_params, pl.logN, B, limb0); }
    return cudaGetLastError();
}

size_t poly_mul_workspace_bytes(const Plan& pl, u32 nl, u32 B) {
    if (pl.logN <= (u32)kMaxLogSlab) return 0;
    return ((size_t)nl * B * sizeof(u64)) << pl.logN;   // one operand-sized scratch
}

cudaError_t launch_poly_mul(const Plan& pl, u32 limb0, u32 nl, const u64* a, const u64* b, u64* c, u64* ws, u32 B, cudaStream_t st) {
    const unsigned nslab = slabs_of(pl, nl, B);
    if (pl.logN <= (u32)kMaxLogSlab) {
        switch (pl.logN) {
#define X(LS) case LS: return launch_row_polymul<LS, true>(pl, limb0, a, b, c, B, nslab, st);
            GFHE_ROW_CASES(X)
#undef X
            default: return cudaErrorInvalidValue;
        }
    }
    if (!ws) return cudaErrorInvalidValue;
    // high stages of a -> c (used as scratch), of b -> ws, in one launch (grid.y picks the operand)
    switch (24 - (int)pl.logN) {
#define X(LC) case LC: { ProfScope ps(pl, K_COL_FWD, st); k_col_fwd<LC><<<dim3(nslab, 2), 256, 0, st>>>(a, c, b, ws, pl.d_fwd, pl.d_params, B, limb0); } break;
        GFHE_COL_CASES(X)
#undef X
        default: return cudaErrorInvalidValue;
    }
    cudaError_t e = launch_row_polymul<12, false>(pl, limb0, c, ws, c, B, nslab, st);
    if (e != cudaSuccess) return e;
    switch (24 - (int)pl.logN) {
#define X(LC) case LC: { ProfScope ps(pl, K_COL_INV, st); k_col_inv<LC><<<nslab, 256, 0, st>>>(c, c, pl.d_inv, pl.d_params, B, limb0); } break;
        GFHE_COL_CASES(X)
#undef X
        default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

}  // namespace gfhe
