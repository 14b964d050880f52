// NTT-domain-resident operations: what an FHE caller does between a forward
// and an inverse transform.  The reference has neither (its NTT is an identity stub, arithmetic.py:15-19); the
// style model is the element-wise family at /root/reference/arithmetic.py:3-13.
//
//   k_mul_acc       out[l, i] = sum_t a[t][l, i] * b[t][l, i]  mod q_l   (key-switching inner product; terms = 1 is the
//                   NTT-domain pointwise product).  Products are accumulated in 128 bits and reduced ONCE, so the kernel
//                   is pure HBM streaming: (16 * terms + 8) bytes per residue, 128-bit coalesced loads.
//   k_automorphism  sigma_k : a(X) -> a(X^k), k odd.  Coefficient domain: a signed index permutation; NTT domain
//                   (bit-reversed evaluation order): the pure permutation out[j] = in[j'] with
//                   2*brv(j') + 1 = k * (2*brv(j) + 1) mod 2N.  Both are gathers with coalesced stores.
#include "plan.h"

namespace gfhe {

typedef unsigned __int128 u128;

__device__ __forceinline__ u64 fold128(u128 z, const LimbParams& P) {
    const u64 z0 = (u64)z, z1 = (u64)(z >> 64);
    const u64 hi = barrett_mul(reduce64(z1, P), P.r64, P);   // z1 * 2^64 mod q
    return csub(hi + reduce64(z0, P), P.q);
}

// grid: (x = chunks of one limb's row, y = limb); each thread owns two adjacent residues per trip (16-byte accesses)
template <int UNROLL>
__global__ void __launch_bounds__(256)
k_mul_acc(const u64* __restrict__ a, const u64* __restrict__ b, u64* __restrict__ out, const LimbParams* __restrict__ params,
          u64 inner, u32 terms, u64 term_stride, u32 limb0) {
    const LimbParams P = params[limb0 + blockIdx.y];
    const u64 row = (u64)blockIdx.y * inner;
    const ulonglong2* a2 = reinterpret_cast<const ulonglong2*>(a + row);
    const ulonglong2* b2 = reinterpret_cast<const ulonglong2*>(b + row);
    ulonglong2* o2 = reinterpret_cast<ulonglong2*>(out + row);
    const u64 n2 = inner >> 1, ts2 = term_stride >> 1;
    const u64 stride = (u64)gridDim.x * blockDim.x;
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n2; i += stride) {
        u128 s0 = 0, s1 = 0;
        u32 t = 0;
        for (; t + UNROLL <= terms; t += UNROLL) {
            ulonglong2 x[UNROLL], y[UNROLL];
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) { x[u] = __ldcs(a2 + i + (u64)(t + u) * ts2); y[u] = __ldcs(b2 + i + (u64)(t + u) * ts2); }
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) { s0 += (u128)x[u].x * y[u].x; s1 += (u128)x[u].y * y[u].y; }
        }
        for (; t < terms; ++t) {
            const ulonglong2 x = __ldcs(a2 + i + (u64)t * ts2), y = __ldcs(b2 + i + (u64)t * ts2);
            s0 += (u128)x.x * y.x; s1 += (u128)x.y * y.y;
        }
        __stcs(o2 + i, make_ulonglong2(fold128(s0, P), fold128(s1, P)));
    }
}

// terms == 1: the NTT-domain pointwise product of canonical residues -- one Barrett product per residue, two
// 16-byte load pairs in flight per thread.
__global__ void __launch_bounds__(256)
k_mul_pointwise(const u64* __restrict__ a, const u64* __restrict__ b, u64* __restrict__ out,
                const LimbParams* __restrict__ params, u64 inner, u32 limb0) {
    const LimbParams P = params[limb0 + blockIdx.y];
    const u64 row = (u64)blockIdx.y * inner;
    const ulonglong2* a2 = reinterpret_cast<const ulonglong2*>(a + row);
    const ulonglong2* b2 = reinterpret_cast<const ulonglong2*>(b + row);
    ulonglong2* o2 = reinterpret_cast<ulonglong2*>(out + row);
    const u64 n2 = inner >> 1;
    const u64 stride = (u64)gridDim.x * blockDim.x;
    u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + stride < n2; i += 2 * stride) {
        const ulonglong2 x0 = __ldcs(a2 + i), y0 = __ldcs(b2 + i), x1 = __ldcs(a2 + i + stride), y1 = __ldcs(b2 + i + stride);
        __stcs(o2 + i, make_ulonglong2(barrett_mul(x0.x, y0.x, P), barrett_mul(x0.y, y0.y, P)));
        __stcs(o2 + i + stride, make_ulonglong2(barrett_mul(x1.x, y1.x, P), barrett_mul(x1.y, y1.y, P)));
    }
    if (i < n2) {
        const ulonglong2 x0 = __ldcs(a2 + i), y0 = __ldcs(b2 + i);
        __stcs(o2 + i, make_ulonglong2(barrett_mul(x0.x, y0.x, P), barrett_mul(x0.y, y0.y, P)));
    }
}

// scalar tail / unaligned variant
__global__ void __launch_bounds__(256)
k_mul_acc_scalar(const u64* __restrict__ a, const u64* __restrict__ b, u64* __restrict__ out,
                 const LimbParams* __restrict__ params, u64 inner, u32 terms, u64 term_stride, u32 limb0) {
    const LimbParams P = params[limb0 + blockIdx.y];
    const u64 row = (u64)blockIdx.y * inner;
    const u64 stride = (u64)gridDim.x * blockDim.x;
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < inner; i += stride) {
        u128 s = 0;
        for (u32 t = 0; t < terms; ++t) s += (u128)a[row + i + (u64)t * term_stride] * b[row + i + (u64)t * term_stride];
        out[row + i] = fold128(s, P);
    }
}

cudaError_t launch_mul_acc(const Plan& pl, u32 limb0, u32 nl, u64 inner, u32 terms, const u64* a, const u64* b,
                           u64 term_stride, u64* out, cudaStream_t st) {
    if (nl == 0 || nl > 65535 || terms == 0) return cudaErrorInvalidValue;
    if (inner == 0) return cudaSuccess;
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, pl.device);
    const bool vec2 = (inner % 2 == 0) && (term_stride % 2 == 0) && ((((uintptr_t)a | (uintptr_t)b | (uintptr_t)out) & 15) == 0);
    const u64 work = vec2 ? inner / 2 : inner;
    u64 per_row = ((u64)sms * 8 + nl - 1) / nl;   // about 8 resident CTAs per SM across all limb rows
    const u64 need = (work + 255) / 256;
    if (per_row > need) per_row = need;
    if (per_row < 1) per_row = 1;
    const dim3 grid((unsigned)per_row, nl);
    ProfScope ps(pl, K_MAC, st);
    if (!vec2) k_mul_acc_scalar<<<grid, 256, 0, st>>>(a, b, out, pl.d_params, inner, terms, term_stride, limb0);
    else if (terms >= 4) k_mul_acc<4><<<grid, 256, 0, st>>>(a, b, out, pl.d_params, inner, terms, term_stride, limb0);
    else if (terms == 1) k_mul_pointwise<<<grid, 256, 0, st>>>(a, b, out, pl.d_params, inner, limb0);
    else k_mul_acc<1><<<grid, 256, 0, st>>>(a, b, out, pl.d_params, inner, terms, term_stride, limb0);
    return cudaGetLastError();
}

// ---- automorphism --------------------------------------------------------------------------------------------------
// grid: (x = polynomial index in [0, nl*B), y = chunk of 1024 outputs inside the polynomial); 4 gathers per thread in flight.
// NTT: kk = k.  Coefficient domain: kk = k^-1 mod 2N (host), out[j] = +-in[(j*kk mod 2N) mod N], minus when j*kk mod 2N >= N.
template <bool NTT>
__global__ void __launch_bounds__(256)
k_automorphism(const u64* __restrict__ in, u64* __restrict__ out, const LimbParams* __restrict__ params, u32 logN, u32 B,
               u32 kk, u32 limb0) {
    const u32 N = 1u << logN, m2 = 2u * N - 1u;
    const u32 poly = blockIdx.x;
    const u64 base = (u64)poly << logN;
    const u64 q = NTT ? 0 : params[limb0 + poly / B].q;
    const u32 j0 = blockIdx.y * 1024u + threadIdx.x;
    u64 v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
        const u32 j = j0 + 256u * u;
        if (j >= N) { v[u] = 0; continue; }
        if (NTT) {
            const u32 e = 2u * (__brev(j) >> (32 - logN)) + 1u;
            const u32 e2 = (e * kk) & m2;                          // odd
            v[u] = __ldg(in + base + (__brev(e2 >> 1) >> (32 - logN)));
        } else {
            const u32 e = (j * kk) & m2;
            const u64 x = __ldg(in + base + (e & (N - 1u)));
            v[u] = (e >= N && x != 0) ? q - x : x;
        }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
        const u32 j = j0 + 256u * u;
        if (j < N) out[base + j] = v[u];
    }
}

cudaError_t launch_automorphism(const Plan& pl, u32 limb0, u32 nl, u32 B, u32 k, bool ntt_domain, const u64* in, u64* out,
                                cudaStream_t st) {
    if (pl.logN == 0 || (k & 1u) == 0) return cudaErrorInvalidValue;
    const u32 N = 1u << pl.logN, m2 = 2u * N - 1u;
    k &= m2;
    u32 kk = k;
    if (!ntt_domain) {   // k^-1 mod 2N by Newton iteration (k odd): x <- x * (2 - k*x), doubling the correct bits
        u32 x = k;       // correct to 3 bits
        for (int i = 0; i < 5; ++i) x *= 2u - k * x;
        kk = x & m2;
    }
    const u64 polys = (u64)nl * B;
    if (polys == 0 || polys > 0x7fffffffull) return cudaErrorInvalidValue;
    const dim3 grid((unsigned)polys, (N + 1023u) / 1024u);
    ProfScope ps(pl, K_AUTO, st);
    if (ntt_domain) k_automorphism<true><<<grid, 256, 0, st>>>(in, out, pl.d_params, pl.logN, B, kk, limb0);
    else k_automorphism<false><<<grid, 256, 0, st>>>(in, out, pl.d_params, pl.logN, B, kk, limb0);
    return cudaGetLastError();
}

}  // namespace gfhe
