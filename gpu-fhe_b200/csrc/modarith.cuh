// 64-bit modular arithmetic on the 32-bit integer datapath of sm_100a (IMAD / IMAD.WIDE /
// IMAD.HI / IADD3).  No tensor cores: the NTT is not a dense contraction.
//
// Replaces the exact-integer semantics of /root/reference/arithmetic.py:3-13 ((a+b)%MOD,
// (a-b)%MOD, (a*b)%MOD on dtype=object arrays) for residues stored as 64-bit words, and
// supplies the butterfly arithmetic the reference's NTT/iNTT stubs (arithmetic.py:15-19)
// never had.
#pragma once
#include <cstddef>
#include <cstdint>

namespace gfhe {

typedef unsigned long long u64;
typedef unsigned int u32;

// Per-limb constants, generated on the host (plan.cu) and resident in HBM.
struct __align__(16) LimbParams {
    u64 q2;         // 2q                      } one 16-byte load (ntt_kernels.inl, limb_consts)
    u64 nq;         // -q mod 2^64             }
    u64 q;          // modulus, odd, < 2^62
    u64 bar_mu;     // floor(2^(k+63) / q), k = bit length of q  (single-mulhi Barrett)
    u64 ninv;       // N^-1 mod q              } 16-byte aligned pair
    u64 ninv_s;     // Shoup companion of ninv }
    u64 w1ninv;     // psi^-brv(1) * N^-1 mod q  (last inverse stage, N^-1 folded in)
    u64 w1ninv_s;
    u64 bar64;      // floor(2^64 / q): one-word Barrett for reducing arbitrary 64-bit inputs
    u32 bar_shift;  // k - 1
    u32 kshift;     // k - 17 (kred: y = x >> kshift fits 32 bits for x < 16q); valid when k >= 20
    u64 kmu;        // floor(2^(32 + kshift) / q) <= 2^16
    u64 r64;        // 2^64 mod q (folds the high word of a 128-bit accumulator, ring_kernels.cu)
    // FP64 policy (q < 2^50): the same constants as doubles, each with its quotient companion c/q
    double qd, qinv;           // q and 1/q
    double ninv_d, ninv_q;     // N^-1 and N^-1/q
    double w1ninv_d, w1ninv_q; // psi^-brv(1) * N^-1 and the same over q
};
static_assert(sizeof(LimbParams) == 144, "LimbParams layout");

// Layout of the warp-tiled twiddle copy (plan.cu fills it, ntt_core.cuh TwT0 reads it).
constexpr u32 kT0Block = 15 * 32;   // entries per warp-tile
__host__ __device__ constexpr size_t t0_entries(size_t N) { return (((N >> 4) + 31) >> 5) * kT0Block; }

// x*w mod q up to a multiple of q: returns x*w - floor(x*ws/2^64)*q in [0, 2q) for ANY 64-bit x,
// given ws = floor(w*2^64/q), w < q, and nq = -q mod 2^64.
// 10 IMAD-class + 1 SEL (+1 zeroing move) in SASS: the exact high product uses
// IMAD.HI / IMAD.WIDE with 64-bit accumulate, the low 64 bits of x*w + hi*nq are one
// IMAD.WIDE chain plus four 32-bit IMADs into the top word.
__device__ __forceinline__ u64 shoup_mul(u64 x, u64 w, u64 ws, u64 nq) {
    u64 r;
    asm("{\n\t"
        ".reg .u32 xl,xh,wl,wh,pl,ph,nl,nh,t,h0,h1,top,lo;\n\t"
        ".reg .u64 acc;\n\t"
        "mov.b64 {xl,xh}, %1;\n\t"
        "mov.b64 {wl,wh}, %2;\n\t"
        "mov.b64 {pl,ph}, %3;\n\t"
        "mov.b64 {nl,nh}, %4;\n\t"
        // (h1:h0) = floor(x * ws / 2^64), exact
        "mul.hi.u32 t, xl, pl;\n\t"
        "mad.lo.cc.u32 t, xh, pl, t;\n\t"
        "madc.hi.u32 h0, xh, pl, 0;\n\t"
        "mad.lo.cc.u32 t, xl, ph, t;\n\t"
        "madc.hi.cc.u32 h0, xl, ph, h0;\n\t"
        "addc.u32 h1, 0, 0;\n\t"
        "mad.lo.cc.u32 h0, xh, ph, h0;\n\t"
        "madc.hi.u32 h1, xh, ph, h1;\n\t"
        // r = x*w + h*nq  (mod 2^64)
        "mul.wide.u32 acc, xl, wl;\n\t"
        "mad.wide.u32 acc, h0, nl, acc;\n\t"
        "mov.b64 {lo,top}, acc;\n\t"
        "mad.lo.u32 top, xl, wh, top;\n\t"
        "mad.lo.u32 top, xh, wl, top;\n\t"
        "mad.lo.u32 top, h0, nh, top;\n\t"
        "mad.lo.u32 top, h1, nl, top;\n\t"
        "mov.b64 %0, {lo,top};\n\t"
        "}\n"
        : "=l"(r)
        : "l"(x), "l"(w), "l"(ws), "l"(nq));
    return r;
}

// x - c if x >= c else x, for x < 2^63 + c and c <= 2^63 (sign of the difference decides).
// 5 SASS instructions: IADD3, IADD3.X, ISETP, SEL, SEL.
__device__ __forceinline__ u64 csub(u64 x, u64 c) {
    u64 d = x - c;
    return (long long)d < 0 ? x : d;
}

// 32-bit form: min(x, x - c) as unsigned -- x - c wraps above x exactly when x < c.  IADD3 + VIMNMX.
__device__ __forceinline__ u32 csub(u32 x, u32 c) { return min(x, x - c); }

__device__ __forceinline__ u64 mulhi64(u64 a, u64 b) { return __umul64hi(a, b); }

// Full reduction of an arbitrary 64-bit value to [0, q) with the one-word Barrett constant.
__device__ __forceinline__ u64 reduce64(u64 x, const LimbParams& P) {
    u64 qh = mulhi64(x, P.bar64);        // floor(x/q) - 1 <= qh <= floor(x/q)
    u64 r = x - qh * P.q;                 // in [0, 2q)
    return csub(r, P.q);
}

// Signed 64-bit storage (torch.int64): Python's (v % q) for possibly negative v.
__device__ __forceinline__ u64 reduce_signed(long long v, const LimbParams& P) {
    if (v >= 0) return reduce64((u64)v, P);
    u64 m = reduce64((u64)(-(v + 1)) + 1ull, P);   // |v| mod q without overflowing at INT64_MIN
    return m == 0 ? 0 : P.q - m;
}

// a*b mod q, fully reduced, for a, b < q < 2^62 (single-mulhi Barrett, quotient error <= 2).
// s = k - 1 (k = bit length of q), mu = floor(2^(k+63) / q).
__device__ __forceinline__ u64 barrett_mul(u64 a, u64 b, u64 q, u64 mu, u32 s) {
    u64 z0 = a * b, z1 = mulhi64(a, b);
    u64 x = (z0 >> s) | (z1 << (64 - s));                  // floor(z / 2^(k-1)) < 2^(k+1)
    u64 qh = mulhi64(x, mu);
    u64 r = z0 - qh * q;                                   // in [0, 3q)
    r = csub(r, 2 * q);
    return csub(r, q);
}
__device__ __forceinline__ u64 barrett_mul(u64 a, u64 b, const LimbParams& P) {
    return barrett_mul(a, b, P.q, P.bar_mu, P.bar_shift);
}

// ---- lazy arithmetic for moduli with four spare bits (2^19 <= q < 2^60) --------------------------
// Approximate-high Shoup product: the quotient estimate xh*ph + hi32(xl*ph) + hi32(xh*pl) drops the
// lo*lo partial product and the low halves of the cross products, so it is the exact one minus 0, 1
// or 2 and the result is
// x*w - h'*q in [0, 4q) for ANY 64-bit x.  5 IMAD.WIDE + 4 IMAD (the exact form needs two IMAD.HI,
// the slowest integer multiply on sm_100a, see tools/pipebench.cu).
__device__ __forceinline__ u64 shoup_a(u64 x, u64 w, u64 ws, u64 nq) {
    u64 r;
    asm("{\n\t"
        ".reg .u32 xl,xh,wl,wh,pl,ph,nl,nh,h0,h1,top,lo,d;\n\t"
        ".reg .u64 t1,t2,h,acc;\n\t"
        "mov.b64 {xl,xh}, %1;\n\t"
        "mov.b64 {wl,wh}, %2;\n\t"
        "mov.b64 {pl,ph}, %3;\n\t"
        "mov.b64 {nl,nh}, %4;\n\t"
        "mul.wide.u32 t1, xl, ph;\n\t"
        "mul.wide.u32 t2, xh, pl;\n\t"
        "mov.b64 {d,h0}, t1;\n\t"
        "mov.b64 {d,h1}, t2;\n\t"
        "add.cc.u32 h0, h0, h1;\n\t"
        "addc.u32 h1, 0, 0;\n\t"
        "mov.b64 t1, {h0,h1};\n\t"
        "mad.wide.u32 h, xh, ph, t1;\n\t"
        "mov.b64 {h0,h1}, h;\n\t"
        "mul.wide.u32 acc, xl, wl;\n\t"
        "mad.wide.u32 acc, h0, nl, acc;\n\t"
        "mov.b64 {lo,top}, acc;\n\t"
        "mad.lo.u32 top, xl, wh, top;\n\t"
        "mad.lo.u32 top, xh, wl, top;\n\t"
        "mad.lo.u32 top, h0, nh, top;\n\t"
        "mad.lo.u32 top, h1, nl, top;\n\t"
        "mov.b64 %0, {lo,top};\n\t"
        "}\n"
        : "=l"(r)
        : "l"(x), "l"(w), "l"(ws), "l"(nq));
    return r;
}

// x mod q up to one q, for x < 16q: k' = floor((x >> kshift) * kmu / 2^32) is floor(x/q) or one less
// (x >> kshift < 2^21, kmu <= 2^16), so x - k'*q is in [0, 2q).  SHF + IMAD.WIDE(.hi) + IMAD.WIDE + IMAD.
__device__ __forceinline__ u64 kred(u64 x, u32 kshift, u32 kmu, u64 nq) {
    const u32 y = (u32)(x >> kshift);
    const u32 k = (u32)(((u64)y * kmu) >> 32);
    return x + (u64)k * nq;
}

// Pointwise product for the Lazy60 pipeline: a, b < 2q  ->  a*b - Q'*q in [0, 6q) (measured < 3.5q), where the
// quotient estimate Q' is the single-mulhi Barrett one taken with the approximate high product (deficit <= 2
// more).  The 128-bit product shares its four partial products between the low and the high word, so the whole
// thing is 8 IMAD.WIDE + 2 IMAD instead of two __umul64hi and two 64-bit multiplies; no final subtractions --
// the inverse butterflies accept [0, 8q).
__device__ __forceinline__ u64 mulhi_a(u64 x, u64 y) {
    const u32 xl = (u32)x, xh = (u32)(x >> 32), yl = (u32)y, yh = (u32)(y >> 32);
    return (u64)xh * yh + (((u64)xl * yh) >> 32) + (((u64)xh * yl) >> 32);
}
__device__ __forceinline__ u64 barrett_lazy(u64 a, u64 b, u64 nq, u64 mu, u32 s) {
    const u32 al = (u32)a, ah = (u32)(a >> 32), bl = (u32)b, bh = (u32)(b >> 32);
    const u64 p0 = (u64)al * bl, p3 = (u64)ah * bh;
    const u64 m1 = (u64)al * bh + (p0 >> 32);          // < 2^64
    const u64 m2 = (u64)ah * bl + (u32)m1;             // < 2^64
    const u64 z0 = (m2 << 32) | (u32)p0;
    const u64 z1 = p3 + (m1 >> 32) + (m2 >> 32);
    const u64 x = (z0 >> s) | (z1 << (64 - s));        // floor(z / 2^(k-1)) < 2^(k+3)
    return z0 + mulhi_a(x, mu) * nq;
}

// 32-bit Shoup product for moduli below 2^30: x*w - floor(x*ws32/2^32)*q in [0, 2q) for ANY 32-bit x, with
// w < q and ws32 = floor(w*2^32/q) -- the high word of the 64-bit companion the tables already hold.
// IMAD.HI + 2 IMAD (the 64-bit product needs 9 multiply instructions).
__device__ __forceinline__ u32 shoup32(u32 x, u32 w, u32 ws32, u32 nq) {
    const u32 h = __umulhi(x, ws32);
    return x * w + h * nq;
}

// Arithmetic policies of the butterfly code (ntt_core.cuh).
//   Harvey  : exact Shoup, any q < 2^62.  forward residues in [0,4q), inverse in [0,2q).
//   Lazy60  : approximate Shoup, 2^19 <= q < 2^60.  forward residues in [0,16q) with one conditional
//             subtraction of 8q every second stage, inverse residues in [0,8q).
//   Shoup32 : q < 2^30.  Harvey's ranges on 32-bit registers: residues are still stored as 64-bit words in HBM
//             (the API's layout), but every butterfly is 3 multiply + 4 add/min instructions instead of 9 + 11.
struct Harvey {
    typedef u64 word;
    typedef u64 store;   // word type of the [L,B,N] arrays in HBM
    typedef ulonglong2 tw_t;    // twiddle-table entry
    static constexpr bool kFloat = false;
    __device__ static __forceinline__ u64 from_u64(u64 v) { return v; }
    __device__ static __forceinline__ u64 to_u64(u64 v) { return v; }
    __device__ static __forceinline__ u64 from_raw(u64 v) { return v; }   // words of the intermediate buffers between passes
    __device__ static __forceinline__ u64 to_raw(u64 v) { return v; }
    static constexpr bool kLazy = false;
    static constexpr int kFwdC = 2;   // forward offset/csub constant = kFwdC * q
    static constexpr int kInvC = 2;
    __device__ static __forceinline__ u64 mul(u64 x, u64 w, u64 ws, u64 nq) { return shoup_mul(x, w, ws, nq); }
    __device__ static __forceinline__ u64 mul_exact(u64 x, u64 w, u64 ws, u64 nq) { return shoup_mul(x, w, ws, nq); }
};
struct Lazy60 {
    typedef u64 word;
    typedef u64 store;   // word type of the [L,B,N] arrays in HBM
    typedef ulonglong2 tw_t;    // twiddle-table entry
    static constexpr bool kFloat = false;
    __device__ static __forceinline__ u64 from_u64(u64 v) { return v; }
    __device__ static __forceinline__ u64 to_u64(u64 v) { return v; }
    __device__ static __forceinline__ u64 from_raw(u64 v) { return v; }   // words of the intermediate buffers between passes
    __device__ static __forceinline__ u64 to_raw(u64 v) { return v; }
    static constexpr bool kLazy = true;
    static constexpr int kFwdC = 4;   // T < 4q
    static constexpr int kInvC = 8;
    __device__ static __forceinline__ u64 mul(u64 x, u64 w, u64 ws, u64 nq) { return shoup_a(x, w, ws, nq); }
    __device__ static __forceinline__ u64 mul_exact(u64 x, u64 w, u64 ws, u64 nq) { return shoup_mul(x, w, ws, nq); }
};
struct Shoup32 {
    typedef u32 word;
    typedef u64 store;   // word type of the [L,B,N] arrays in HBM
    typedef u64 tw_t;    // twiddle-table entry (compact: w | w' << 32)
    static constexpr bool kFloat = false;
    __device__ static __forceinline__ u32 from_u64(u64 v) { return (u32)v; }
    __device__ static __forceinline__ u64 to_u64(u32 v) { return v; }
    __device__ static __forceinline__ u32 from_raw(u64 v) { return (u32)v; }
    __device__ static __forceinline__ u64 to_raw(u32 v) { return v; }
    static constexpr bool kLazy = false;
    static constexpr int kFwdC = 2;
    static constexpr int kInvC = 2;
    __device__ static __forceinline__ u32 mul(u32 x, u64 w, u64 ws, u32 nq) { return shoup32(x, (u32)w, (u32)(ws >> 32), nq); }
    __device__ static __forceinline__ u32 mul_exact(u32 x, u64 w, u64 ws, u32 nq) { return shoup32(x, (u32)w, (u32)(ws >> 32), nq); }
};

// Shoup32 on arrays of 32-bit words in HBM (int32 / uint32 tensors): half the traffic of the 64-bit layout.
struct Shoup32P : Shoup32 {
    typedef u32 store;
};

// ---- FP64 arithmetic for moduli below 2^50 -----------------------------------------------------------------------
// B200 retires DFMA at the same rate as IMAD, on a pipe the integer kernels leave idle, and one DFMA pair gives the
// exact 106-bit product of two 53-bit integers (h = fl(y*w), l = fma(y, w, -h)).  Residues are integer-valued doubles
// kept in (-2^51, 2^51); twiddles are stored as {w, fl(w/q)} (bit patterns in the same ulonglong2 tables).
//   fp_mulmod : y*w - rint(y*w/q)*q, exact, in [-0.75q, 0.75q] for |y| < 2^51   (6 FP64 instructions)
//   fp_reduce : x - rint(x/q)*q, exact, in [-0.51q, 0.51q]                        (3 FP64 instructions)
constexpr double kFpMagic = 6755399441055744.0;   // 1.5 * 2^52: fma(v, 1, M) - M = rint(v) for |v| < 2^51
__device__ __forceinline__ double fp_mulmod(double y, double w, double wq, double q) {
    const double c = fma(y, wq, kFpMagic) - kFpMagic;
    const double h = y * w;
    const double l = fma(y, w, -h);
    return fma(-c, q, h) + l;
}
__device__ __forceinline__ double fp_reduce(double x, double q, double qinv) {
    const double c = fma(x, qinv, kFpMagic) - kFpMagic;
    return fma(-c, q, x);
}
__device__ __forceinline__ double fp_canon(double x, double q) { return x < 0.0 ? x + q : x; }   // [-q, q) -> [0, q)
__device__ __forceinline__ double bits_to_double(u64 b) { return __longlong_as_double((long long)b); }

struct Fp50 {
    typedef double word;
    typedef u64 store;   // word type of the [L,B,N] arrays in HBM
    typedef ulonglong2 tw_t;    // twiddle-table entry
    static constexpr bool kFloat = true;
    static constexpr bool kLazy = false;
    static constexpr int kFwdC = 2;
    static constexpr int kInvC = 2;
    // v < 2^52: place v in the mantissa of 2^52 and subtract 2^52 (LOP3 + DADD, no conversion instruction)
    __device__ static __forceinline__ double from_u64(u64 v) {
        return __longlong_as_double((long long)(v | 0x4330000000000000ull)) - 4503599627370496.0;
    }
    // 0 <= v < 2^52, integer-valued
    __device__ static __forceinline__ u64 to_u64(double v) {
        return (u64)__double_as_longlong(v + 4503599627370496.0) & 0x000FFFFFFFFFFFFFull;
    }
    // between the two passes of a large transform the (signed, unreduced) doubles travel as their bit patterns
    __device__ static __forceinline__ double from_raw(u64 v) { return __longlong_as_double((long long)v); }
    __device__ static __forceinline__ u64 to_raw(double v) { return (u64)__double_as_longlong(v); }
};

}  // namespace gfhe
