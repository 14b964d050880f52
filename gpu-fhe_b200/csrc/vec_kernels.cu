// Element-wise modular add / sub / mul over [Lq, inner] with one modulus per leading index.
// Exact replacements for /root/reference/arithmetic.py:3-13 ((a+b)%MOD, (a-b)%MOD, (a*b)%MOD
// in Python-int semantics): inputs are int64 words of ANY value (negative included), outputs
// are canonical residues in [0, MOD).  Pure HBM streaming: 24 bytes per residue, 128-bit
// coalesced accesses, grid sized in multiples of the SM count.
#include "plan.h"

namespace gfhe {

// UNS: the words are unsigned 64-bit values (uint64 containers; values >= 2^63 are NOT negative numbers)
template <int OP, bool UNS>
__device__ __forceinline__ long long vec_elem(long long a, long long b, const LimbParams& P) {
    const u64 x = UNS ? reduce64((u64)a, P) : reduce_signed(a, P), y = UNS ? reduce64((u64)b, P) : reduce_signed(b, P);
    if (OP == 0) return (long long)csub(x + y, P.q);
    if (OP == 1) return (long long)(x >= y ? x - y : x + P.q - y);
    return (long long)barrett_mul(x, y, P);
}

// grid: (x = chunks of one modulus' row, y = modulus index)
template <int OP, bool VEC2, bool UNS>
__global__ void __launch_bounds__(256)
k_vec(const long long* __restrict__ a, const long long* __restrict__ b, long long* __restrict__ out,
      const LimbParams* __restrict__ params, u64 inner, u32 limb0) {
    const LimbParams P = params[limb0 + blockIdx.y];
    const u64 row = (u64)blockIdx.y * inner;
    const u64 stride = (u64)gridDim.x * blockDim.x;
    u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (VEC2) {
        const longlong2* a2 = reinterpret_cast<const longlong2*>(a + row);
        const longlong2* b2 = reinterpret_cast<const longlong2*>(b + row);
        longlong2* o2 = reinterpret_cast<longlong2*>(out + row);
        const u64 n2 = inner >> 1;
        for (; i + stride < n2; i += 2 * stride) {   // two 16-byte load pairs in flight per thread, streaming cache hints
            const longlong2 x0 = __ldcs(a2 + i), y0 = __ldcs(b2 + i), x1 = __ldcs(a2 + i + stride), y1 = __ldcs(b2 + i + stride);
            longlong2 z0, z1;
            z0.x = vec_elem<OP, UNS>(x0.x, y0.x, P); z0.y = vec_elem<OP, UNS>(x0.y, y0.y, P);
            z1.x = vec_elem<OP, UNS>(x1.x, y1.x, P); z1.y = vec_elem<OP, UNS>(x1.y, y1.y, P);
            __stcs(o2 + i, z0);
            __stcs(o2 + i + stride, z1);
        }
        if (i < n2) {
            const longlong2 x = __ldcs(a2 + i), y = __ldcs(b2 + i);
            longlong2 z;
            z.x = vec_elem<OP, UNS>(x.x, y.x, P);
            z.y = vec_elem<OP, UNS>(x.y, y.y, P);
            __stcs(o2 + i, z);
        }
    } else {
        for (; i < inner; i += stride) out[row + i] = vec_elem<OP, UNS>(a[row + i], b[row + i], P);
    }
}

template <int OP, bool UNS>
static cudaError_t launch_vec(const Plan& pl, u32 limb0, u32 nl, const long long* a, const long long* b, long long* out,
                              u64 inner, cudaStream_t st) {
    if (inner == 0) return cudaSuccess;
    if (nl > 65535u) {   // grid.y limit: one launch per 65535 moduli (a MOD that varies along a long axis, e.g. per coefficient)
        for (u32 l = 0; l < nl; l += 65535u) {
            const u32 n = nl - l < 65535u ? nl - l : 65535u;
            if (cudaError_t e = launch_vec<OP, UNS>(pl, limb0 + l, n, a + (u64)l * inner, b + (u64)l * inner, out + (u64)l * inner, inner, st)) return e;
        }
        return cudaSuccess;
    }
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, pl.device);
    const bool vec2 = (inner % 2 == 0) && ((((uintptr_t)a | (uintptr_t)b | (uintptr_t)out) & 15) == 0);
    const u64 work = vec2 ? inner / 2 : inner;
    // about 8 resident CTAs per SM across all modulus rows, never more CTAs than work
    u64 per_row = ((u64)sms * 8 + nl - 1) / nl;
    const u64 need = (work + 255) / 256;
    if (per_row > need) per_row = need;
    if (per_row < 1) per_row = 1;
    dim3 grid((unsigned)per_row, nl);
    ProfScope ps(pl, K_VEC, st);
    if (vec2) k_vec<OP, true, UNS><<<grid, 256, 0, st>>>(a, b, out, pl.d_params, inner, limb0);
    else      k_vec<OP, false, UNS><<<grid, 256, 0, st>>>(a, b, out, pl.d_params, inner, limb0);
    return cudaGetLastError();
}

// op: 0 add, 1 sub, 2 mul; + 4 when the words are unsigned
cudaError_t launch_vec_op(int op, const Plan& pl, u32 limb0, u32 nl, const long long* a, const long long* b,
                          long long* out, u64 inner, cudaStream_t st) {
    if (nl == 0) return cudaErrorInvalidValue;
    switch (op) {
        case 0: return launch_vec<0, false>(pl, limb0, nl, a, b, out, inner, st);
        case 1: return launch_vec<1, false>(pl, limb0, nl, a, b, out, inner, st);
        case 2: return launch_vec<2, false>(pl, limb0, nl, a, b, out, inner, st);
        case 4: return launch_vec<0, true>(pl, limb0, nl, a, b, out, inner, st);
        case 5: return launch_vec<1, true>(pl, limb0, nl, a, b, out, inner, st);
        case 6: return launch_vec<2, true>(pl, limb0, nl, a, b, out, inner, st);
        default: return cudaErrorInvalidValue;
    }
}

}  // namespace gfhe
