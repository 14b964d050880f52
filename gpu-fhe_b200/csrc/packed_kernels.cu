// Packed 32-bit layout (int32 / uint32 arrays, every modulus of the plan below 2^30): the operations around the
// transforms -- element-wise add / sub / mul (/root/reference/arithmetic.py:3-13 semantics: ANY int32 inputs, canonical
// outputs), the NTT-domain multiply-accumulate and the Galois automorphism.  Same contracts as vec_kernels.cu and
// ring_kernels.cu, half the bytes: 12 B per residue for the element-wise ops, (8 * terms + 4) for the inner product.
// All pure HBM streaming with 16-byte accesses (four residues per thread and trip).
#include "plan.h"

namespace gfhe {

// 32-bit arithmetic throughout (the 64-bit helpers of vec_kernels.cu make these kernels issue-bound at 12 B per residue).
struct Q32 {
    u32 q, bar32, c31; u64 bar64;   // floor(2^32 / q), 2^31 mod q, floor(2^64 / q)
    __device__ explicit Q32(const LimbParams& P) : q((u32)P.q), bar32((u32)(P.bar64 >> 32)), c31(0x80000000u % (u32)P.q), bar64(P.bar64) {}
    // any int32 -> [0, q): a + 2^31 is a u32; one-word Barrett (quotient short by at most 1), then take 2^31 back off
    __device__ __forceinline__ u32 canon(int a) const {
        const u32 u = (u32)a ^ 0x80000000u;
        const u32 r = csub(u - __umulhi(u, bar32) * q, q);
        return r >= c31 ? r - c31 : r + q - c31;
    }
    // any u64 -> [0, q): floor(s/q) - 1 <= hi(s * bar64) <= floor(s/q), so s - hi*q < 2q fits the low word
    __device__ __forceinline__ u32 red64(u64 s) const { return csub((u32)s - (u32)__umul64hi(s, bar64) * q, q); }
};

template <int OP>
__device__ __forceinline__ int vec_elem32(int a, int b, const Q32& Q) {
    const u32 x = Q.canon(a), y = Q.canon(b);
    if (OP == 0) return (int)csub(x + y, Q.q);
    if (OP == 1) return (int)(x >= y ? x - y : x + Q.q - y);
    return (int)Q.red64((u64)x * y);
}

// grid: (x = chunks of one modulus' row, y = modulus index)
template <int OP, bool VEC4>
__global__ void __launch_bounds__(256)
k_vec32(const int* __restrict__ a, const int* __restrict__ b, int* __restrict__ out, const LimbParams* __restrict__ params,
        u64 inner, u32 limb0) {
    const Q32 P(params[limb0 + blockIdx.y]);
    const u64 row = (u64)blockIdx.y * inner;
    const u64 stride = (u64)gridDim.x * blockDim.x;
    u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (VEC4) {
        const int4* a4 = reinterpret_cast<const int4*>(a + row);
        const int4* b4 = reinterpret_cast<const int4*>(b + row);
        int4* o4 = reinterpret_cast<int4*>(out + row);
        const u64 n4 = inner >> 2;
        for (; i + stride < n4; i += 2 * stride) {   // two 16-byte load pairs in flight per thread
            const int4 x0 = __ldcs(a4 + i), y0 = __ldcs(b4 + i), x1 = __ldcs(a4 + i + stride), y1 = __ldcs(b4 + i + stride);
            __stcs(o4 + i, make_int4(vec_elem32<OP>(x0.x, y0.x, P), vec_elem32<OP>(x0.y, y0.y, P), vec_elem32<OP>(x0.z, y0.z, P),
                                     vec_elem32<OP>(x0.w, y0.w, P)));
            __stcs(o4 + i + stride, make_int4(vec_elem32<OP>(x1.x, y1.x, P), vec_elem32<OP>(x1.y, y1.y, P),
                                              vec_elem32<OP>(x1.z, y1.z, P), vec_elem32<OP>(x1.w, y1.w, P)));
        }
        if (i < n4) {
            const int4 x0 = __ldcs(a4 + i), y0 = __ldcs(b4 + i);
            __stcs(o4 + i, make_int4(vec_elem32<OP>(x0.x, y0.x, P), vec_elem32<OP>(x0.y, y0.y, P), vec_elem32<OP>(x0.z, y0.z, P),
                                     vec_elem32<OP>(x0.w, y0.w, P)));
        }
    } else {
        for (; i < inner; i += stride) out[row + i] = vec_elem32<OP>(a[row + i], b[row + i], P);
    }
}

static dim3 stream_grid(const Plan& pl, u32 nl, u64 work) {   // about 8 resident CTAs per SM across all rows, never more than the work
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, pl.device);
    u64 per_row = ((u64)sms * 8 + nl - 1) / nl;
    const u64 need = (work + 255) / 256;
    if (per_row > need) per_row = need;
    if (per_row < 1) per_row = 1;
    return dim3((unsigned)per_row, nl);
}

template <int OP>
static cudaError_t launch_vec32(const Plan& pl, u32 limb0, u32 nl, const int* a, const int* b, int* out, u64 inner, cudaStream_t st) {
    const bool vec4 = (inner % 4 == 0) && ((((uintptr_t)a | (uintptr_t)b | (uintptr_t)out) & 15) == 0);
    const dim3 grid = stream_grid(pl, nl, vec4 ? inner / 4 : inner);
    ProfScope ps(pl, K_VEC, st);
    if (vec4) k_vec32<OP, true><<<grid, 256, 0, st>>>(a, b, out, pl.d_params, inner, limb0);
    else      k_vec32<OP, false><<<grid, 256, 0, st>>>(a, b, out, pl.d_params, inner, limb0);
    return cudaGetLastError();
}

cudaError_t launch_vec_op_u32(int op, const Plan& pl, u32 limb0, u32 nl, const int* a, const int* b, int* out, u64 inner,
                              cudaStream_t st) {
    if (nl == 0 || !pl.w32) return cudaErrorInvalidValue;
    if (inner == 0) return cudaSuccess;
    if (nl > 65535u) {   // grid.y limit: one launch per 65535 moduli
        for (u32 l = 0; l < nl; l += 65535u) {
            const u32 n = nl - l < 65535u ? nl - l : 65535u;
            if (cudaError_t e = launch_vec_op_u32(op, pl, limb0 + l, n, a + (u64)l * inner, b + (u64)l * inner, out + (u64)l * inner, inner, st)) return e;
        }
        return cudaSuccess;
    }
    switch (op) {
        case 0: return launch_vec32<0>(pl, limb0, nl, a, b, out, inner, st);
        case 1: return launch_vec32<1>(pl, limb0, nl, a, b, out, inner, st);
        case 2: return launch_vec32<2>(pl, limb0, nl, a, b, out, inner, st);
        default: return cudaErrorInvalidValue;
    }
}

// ---- NTT-domain multiply-accumulate: out[l, i] = sum_t a[t][l, i] * b[t][l, i] mod q_l, canonical operands ------------
// Products are below 2^60, so 8 of them fit a 64-bit accumulator; it is folded below q every 8 terms.
template <bool VEC4>
__global__ void __launch_bounds__(256)
k_mul_acc32(const u32* __restrict__ a, const u32* __restrict__ b, u32* __restrict__ out, const LimbParams* __restrict__ params,
            u64 inner, u32 terms, u64 term_stride, u32 limb0) {
    const Q32 Q(params[limb0 + blockIdx.y]);
    const u64 row = (u64)blockIdx.y * inner;
    const u64 stride = (u64)gridDim.x * blockDim.x;
    if (VEC4) {
        const uint4* a4 = reinterpret_cast<const uint4*>(a + row);
        const uint4* b4 = reinterpret_cast<const uint4*>(b + row);
        uint4* o4 = reinterpret_cast<uint4*>(out + row);
        const u64 n4 = inner >> 2, ts4 = term_stride >> 2;
        for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
            u64 s0 = 0, s1 = 0, s2 = 0, s3 = 0;
            for (u32 t0 = 0; t0 < terms; t0 += 8) {
                const u32 te = t0 + 8 < terms ? t0 + 8 : terms;
#pragma unroll 4
                for (u32 t = t0; t < te; ++t) {
                    const uint4 x = __ldcs(a4 + i + (u64)t * ts4), y = __ldcs(b4 + i + (u64)t * ts4);
                    s0 += (u64)x.x * y.x; s1 += (u64)x.y * y.y; s2 += (u64)x.z * y.z; s3 += (u64)x.w * y.w;
                }
                s0 = Q.red64(s0); s1 = Q.red64(s1); s2 = Q.red64(s2); s3 = Q.red64(s3);   // < q < 2^30
            }
            __stcs(o4 + i, make_uint4((u32)s0, (u32)s1, (u32)s2, (u32)s3));
        }
    } else {
        for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < inner; i += stride) {
            u64 s = 0;
            for (u32 t = 0; t < terms; ++t) {
                s += (u64)a[row + i + (u64)t * term_stride] * b[row + i + (u64)t * term_stride];
                if ((t & 7u) == 7u) s = Q.red64(s);
            }
            out[row + i] = Q.red64(s);
        }
    }
}

cudaError_t launch_mul_acc_u32(const Plan& pl, u32 limb0, u32 nl, u64 inner, u32 terms, const u32* a, const u32* b,
                               u64 term_stride, u32* out, cudaStream_t st) {
    if (nl == 0 || nl > 65535 || terms == 0 || !pl.w32) return cudaErrorInvalidValue;
    if (inner == 0) return cudaSuccess;
    const bool vec4 = (inner % 4 == 0) && (term_stride % 4 == 0) && ((((uintptr_t)a | (uintptr_t)b | (uintptr_t)out) & 15) == 0);
    const dim3 grid = stream_grid(pl, nl, vec4 ? inner / 4 : inner);
    ProfScope ps(pl, K_MAC, st);
    // one term = the pointwise product: the element-wise kernel keeps two load pairs in flight (5.9 vs 5.3 TB/s)
    if (vec4 && terms == 1) k_vec32<2, true><<<grid, 256, 0, st>>>(reinterpret_cast<const int*>(a), reinterpret_cast<const int*>(b),
                                                                  reinterpret_cast<int*>(out), pl.d_params, inner, limb0);
    else if (vec4) k_mul_acc32<true><<<grid, 256, 0, st>>>(a, b, out, pl.d_params, inner, terms, term_stride, limb0);
    else      k_mul_acc32<false><<<grid, 256, 0, st>>>(a, b, out, pl.d_params, inner, terms, term_stride, limb0);
    return cudaGetLastError();
}

// ---- automorphism a(X) -> a(X^k): the index maps of ring_kernels.cu on 32-bit words ------------------------------------
template <bool NTT>
__global__ void __launch_bounds__(256)
k_automorphism32(const u32* __restrict__ in, u32* __restrict__ out, const LimbParams* __restrict__ params, u32 logN, u32 B,
                 u32 kk, u32 limb0) {
    const u32 N = 1u << logN, m2 = 2u * N - 1u;
    const u32 poly = blockIdx.x;
    const u64 base = (u64)poly << logN;
    const u32 q = NTT ? 0u : (u32)params[limb0 + poly / B].q;
    const u32 j0 = blockIdx.y * 1024u + threadIdx.x;
    u32 v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
        const u32 j = j0 + 256u * u;
        if (j >= N) { v[u] = 0; continue; }
        if (NTT) {
            const u32 e = 2u * (__brev(j) >> (32 - logN)) + 1u;
            const u32 e2 = (e * kk) & m2;
            v[u] = __ldg(in + base + (__brev(e2 >> 1) >> (32 - logN)));
        } else {
            const u32 e = (j * kk) & m2;
            const u32 x = __ldg(in + base + (e & (N - 1u)));
            v[u] = (e >= N && x != 0) ? q - x : x;
        }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
        const u32 j = j0 + 256u * u;
        if (j < N) out[base + j] = v[u];
    }
}

cudaError_t launch_automorphism_u32(const Plan& pl, u32 limb0, u32 nl, u32 B, u32 k, bool ntt_domain, const u32* in, u32* out,
                                    cudaStream_t st) {
    if (pl.logN == 0 || (k & 1u) == 0 || !pl.w32) return cudaErrorInvalidValue;
    const u32 N = 1u << pl.logN, m2 = 2u * N - 1u;
    k &= m2;
    u32 kk = k;
    if (!ntt_domain) {   // k^-1 mod 2N by Newton iteration, as in ring_kernels.cu
        u32 x = k;
        for (int i = 0; i < 5; ++i) x *= 2u - k * x;
        kk = x & m2;
    }
    const u64 polys = (u64)nl * B;
    if (polys == 0 || polys > 0x7fffffffull) return cudaErrorInvalidValue;
    const dim3 grid((unsigned)polys, (N + 1023u) / 1024u);
    ProfScope ps(pl, K_AUTO, st);
    if (ntt_domain) k_automorphism32<true><<<grid, 256, 0, st>>>(in, out, pl.d_params, pl.logN, B, kk, limb0);
    else k_automorphism32<false><<<grid, 256, 0, st>>>(in, out, pl.d_params, pl.logN, B, kk, limb0);
    return cudaGetLastError();
}

}  // namespace gfhe
