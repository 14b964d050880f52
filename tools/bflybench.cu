// Butterfly-variant microbenchmark for sm_100a: issue-clocks per 60-bit butterfly for candidate
// arithmetic (exact vs approximate high product, exact vs high-word conditional subtraction).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o bflybench tools/bflybench.cu && ./bflybench
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>
#include "../gpu-fhe_b200/csrc/modarith.cuh"
using namespace gfhe;

constexpr int ITERS = 512, CH = 8;

// high-word conditional subtraction: subtracts c when hi32(x) > hi32(c); result < max(c + 2^32, x - c)
__device__ __forceinline__ u64 csub_hi(u64 x, u64 c) {
    u64 r;
    asm("{\n\t"
        ".reg .u32 xl,xh,cl,ch;\n\t"
        ".reg .pred p;\n\t"
        "mov.b64 {xl,xh}, %1;\n\t"
        "mov.b64 {cl,ch}, %2;\n\t"
        "setp.gt.u32 p, xh, ch;\n\t"
        "@p sub.cc.u32 xl, xl, cl;\n\t"
        "@p subc.u32 xh, xh, ch;\n\t"
        "mov.b64 %0, {xl,xh};\n\t"
        "}\n" : "=l"(r) : "l"(x), "l"(c));
    return r;
}

// FP64-assisted product: x*w mod q from the split x = xh*2^32 + xl with w0 = w, w1 = w*2^32 mod q and
// the quotient of each half estimated on the FP64 pipe (r0 = w0/q, r1 = w1/q as doubles).
// Returns r + K (mod 2^64), r in [-1.01q, 1.01q], K = 0x43300000 * nql * 2^32 a per-limb constant.
__device__ __forceinline__ u64 dfmul(u64 x, u64 w0, u64 w1, double r0, double r1, u64 nq) {
    const u32 xl = (u32)x, xh = (u32)(x >> 32);
    const double ul = __hiloint2double(0x43300000, (int)xl) - 4503599627370496.0;
    const double uh = __hiloint2double(0x43300000, (int)xh) - 4503599627370496.0;
    const double R = fma(ul, r0, fma(uh, r1, 4503599627370496.0));
    const u32 hlo = (u32)__double2loint(R), hhi = (u32)__double2hiint(R);
    u64 r;
    asm("{\n\t"
        ".reg .u32 al,ah,bl,bh,nl,nh,top,lo;\n\t"
        ".reg .u64 acc;\n\t"
        "mov.b64 {al,ah}, %3;\n\t"
        "mov.b64 {bl,bh}, %4;\n\t"
        "mov.b64 {nl,nh}, %5;\n\t"
        "mul.wide.u32 acc, %1, al;\n\t"
        "mad.wide.u32 acc, %2, bl, acc;\n\t"
        "mad.wide.u32 acc, %6, nl, acc;\n\t"
        "mov.b64 {lo,top}, acc;\n\t"
        "mad.lo.u32 top, %1, ah, top;\n\t"
        "mad.lo.u32 top, %2, bh, top;\n\t"
        "mad.lo.u32 top, %6, nh, top;\n\t"
        "mad.lo.u32 top, %7, nl, top;\n\t"
        "mov.b64 %0, {lo,top};\n\t"
        "}\n" : "=l"(r) : "r"(xl), "r"(xh), "l"(w0), "l"(w1), "l"(nq), "r"(hlo), "r"(hhi));
    return r;
}

// same, but the two 32-bit halves are converted with I2F.F64.U32 (conversion pipe) instead of the
// 2^52 bit trick (2 moves + 2 DADD)
__device__ __forceinline__ u64 dfmul_cvt(u64 x, u64 w0, u64 w1, double r0, double r1, u64 nq) {
    const u32 xl = (u32)x, xh = (u32)(x >> 32);
    const double ul = __uint2double_rn(xl), uh = __uint2double_rn(xh);
    const double R = fma(ul, r0, fma(uh, r1, 4503599627370496.0));
    const u32 hlo = (u32)__double2loint(R), hhi = (u32)__double2hiint(R);
    u64 r;
    asm("{\n\t"
        ".reg .u32 al,ah,bl,bh,nl,nh,top,lo;\n\t"
        ".reg .u64 acc;\n\t"
        "mov.b64 {al,ah}, %3;\n\t"
        "mov.b64 {bl,bh}, %4;\n\t"
        "mov.b64 {nl,nh}, %5;\n\t"
        "mul.wide.u32 acc, %1, al;\n\t"
        "mad.wide.u32 acc, %2, bl, acc;\n\t"
        "mad.wide.u32 acc, %6, nl, acc;\n\t"
        "mov.b64 {lo,top}, acc;\n\t"
        "mad.lo.u32 top, %1, ah, top;\n\t"
        "mad.lo.u32 top, %2, bh, top;\n\t"
        "mad.lo.u32 top, %6, nh, top;\n\t"
        "mad.lo.u32 top, %7, nl, top;\n\t"
        "mov.b64 %0, {lo,top};\n\t"
        "}\n" : "=l"(r) : "r"(xl), "r"(xh), "l"(w0), "l"(w1), "l"(nq), "r"(hlo), "r"(hhi));
    return r;
}


// ---- split representation: a residue is two 32-bit halves, each the LOW word of a register pair whose HIGH word
// holds 0x43300000, i.e. the pair read as a double is 2^52 + half.  Integer adds write the low words directly
// (32-bit IADD3 / IADD3.X can target any register), so the int->double conversion costs one DADD and no moves.
struct Split { double l, h; };
__device__ __forceinline__ Split split_make(u64 x) {
    Split s; s.l = __hiloint2double(0x43300000, (int)(u32)x); s.h = __hiloint2double(0x43300000, (int)(u32)(x >> 32)); return s;
}
__device__ __forceinline__ u64 split_join(const Split& s) { return ((u64)(u32)__double2hiint(s.h) * 0 + ((u64)(u32)__double2loint(s.h) << 32)) | (u32)__double2loint(s.l); }
// T = y*w - h*q (+K) from the split y; returns 64-bit integer
template <bool NOFP = false>
__device__ __forceinline__ u64 dfmul_split(const Split& y, u64 w0, u64 w1, double r0, double r1, u64 nq) {
    u32 hlo, hhi;
    if (NOFP) { hlo = (u32)__double2loint(y.l) ^ (u32)__double2loint(y.h); hhi = 0x43300001u; }
    else {
    const double ul = y.l - 4503599627370496.0, uh = y.h - 4503599627370496.0;
    const double R = fma(ul, r0, fma(uh, r1, 4503599627370496.0));
    hlo = (u32)__double2loint(R); hhi = (u32)__double2hiint(R); }
    const u32 yl = (u32)__double2loint(y.l), yh = (u32)__double2loint(y.h);
    u64 r;
    asm("{\n\t"
        ".reg .u32 al,ah,bl,bh,nl,nh,top,lo;\n\t"
        ".reg .u64 acc;\n\t"
        "mov.b64 {al,ah}, %3;\n\t"
        "mov.b64 {bl,bh}, %4;\n\t"
        "mov.b64 {nl,nh}, %5;\n\t"
        "mul.wide.u32 acc, %1, al;\n\t"
        "mad.wide.u32 acc, %2, bl, acc;\n\t"
        "mad.wide.u32 acc, %6, nl, acc;\n\t"
        "mov.b64 {lo,top}, acc;\n\t"
        "mad.lo.u32 top, %1, ah, top;\n\t"
        "mad.lo.u32 top, %2, bh, top;\n\t"
        "mad.lo.u32 top, %6, nh, top;\n\t"
        "mad.lo.u32 top, %7, nl, top;\n\t"
        "mov.b64 %0, {lo,top};\n\t"
        "}\n" : "=l"(r) : "r"(yl), "r"(yh), "l"(w0), "l"(w1), "l"(nq), "r"(hlo), "r"(hhi));
    return r;
}
// ---- round-2 candidates -------------------------------------------------------------------------------------------
// Montgomery-friendly modulus q = 1 (mod 2^32): -q mod 2^64 = nh*2^32 + (2^32 - 1), so h*(-q) = h*(nh+1)*2^32 - h and the
// IMAD.WIDE h0*nl plus the IMAD h1*nl of shoup_a become a 64-bit subtraction and a 32-bit one.  4 IMAD.WIDE + 3 IMAD.
__device__ __forceinline__ u64 shoup_mf(u64 x, u64 w, u64 ws, u32 nhp1) {
    u64 r;
    asm("{\n\t"
        ".reg .u32 xl,xh,wl,wh,pl,ph,h0,h1,top,lo,d;\n\t"
        ".reg .u64 t1,t2,h,acc;\n\t"
        "mov.b64 {xl,xh}, %1;\n\t"
        "mov.b64 {wl,wh}, %2;\n\t"
        "mov.b64 {pl,ph}, %3;\n\t"
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
        "mov.b64 {lo,top}, acc;\n\t"
        "sub.cc.u32 lo, lo, h0;\n\t"
        "subc.u32 top, top, h1;\n\t"       // borrow and -h1 in one instruction
        "mad.lo.u32 top, xl, wh, top;\n\t"
        "mad.lo.u32 top, xh, wl, top;\n\t"
        "mad.lo.u32 top, h0, %4, top;\n\t"
        "mov.b64 %0, {lo,top};\n\t"
        "}\n"
        : "=l"(r)
        : "l"(x), "l"(w), "l"(ws), "r"(nhp1));
    return r;
}
// SURVEY 8d family q = 2^60 - c, c < 2^32: the high word of -q is 0xF0000000, so h0*nh is a shift.  5 IMAD.WIDE + 3 IMAD + shift-add.
__device__ __forceinline__ u64 shoup_sol(u64 x, u64 w, u64 ws, u64 nq) {
    u64 r;
    asm("{\n\t"
        ".reg .u32 xl,xh,wl,wh,pl,ph,nl,nh,h0,h1,top,lo,d,s;\n\t"
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
        "shl.b32 s, h0, 28;\n\t"
        "sub.u32 top, top, s;\n\t"          // h0 * 0xF0000000 = -(h0 << 28) mod 2^32
        "mad.lo.u32 top, h1, nl, top;\n\t"
        "mov.b64 %0, {lo,top};\n\t"
        "}\n"
        : "=l"(r)
        : "l"(x), "l"(w), "l"(ws), "l"(nq));
    return r;
}
// Montgomery product of two general residues (pointwise product with R^-1 left for the inverse's constants): a*b*2^-64 mod q
// in [0, 2q) for a*b < q*2^64.  qinv = -q^-1 mod 2^64.  The low halves cancel exactly, so the carry is (z0 != 0).
__device__ __forceinline__ u64 mont_mul(u64 a, u64 b, u64 q, u64 qinv) {
    const u64 z0 = a * b, z1 = __umul64hi(a, b);
    const u64 m = z0 * qinv;
    return z1 + __umul64hi(m, q) + (z0 != 0);
}

// s = (a op b) + c on split halves, three-input 64-bit add with the result written into the low words of s's pairs
__device__ __forceinline__ void split_add3(Split& s, u32 al, u32 ah, u64 t, u64 c, bool sub) {
    u32 lo, hi;
    const u32 tl = (u32)t, th = (u32)(t >> 32), cl = (u32)c, ch = (u32)(c >> 32);
    if (!sub) asm("{.reg .u32 x; add.cc.u32 x, %2, %4; addc.cc.u32 %1, %3, %5; add.cc.u32 %0, x, %6; addc.u32 %1, %1, %7;}"
                  : "=r"(lo), "=&r"(hi) : "r"(al), "r"(ah), "r"(tl), "r"(th), "r"(cl), "r"(ch));
    else asm("{.reg .u32 x; sub.cc.u32 x, %2, %4; subc.cc.u32 %1, %3, %5; add.cc.u32 %0, x, %6; addc.u32 %1, %1, %7;}"
             : "=r"(lo), "=&r"(hi) : "r"(al), "r"(ah), "r"(tl), "r"(th), "r"(cl), "r"(ch));
    s.l = __hiloint2double(__double2hiint(s.l), (int)lo);
    s.h = __hiloint2double(__double2hiint(s.h), (int)hi);
}

template <int MODE>
__global__ void __launch_bounds__(512, 2) kb(u64* out, long long* cyc, u64 q, u64 w, u64 ws) {
    u64 X[CH], Y[CH];
    const u64 q2 = 2 * q, q4 = 4 * q, nq = 0 - q;
    const double r0 = (double)w / (double)q, r1 = (double)ws / 18446744073709551616.0;
#pragma unroll
    for (int i = 0; i < CH; ++i) { X[i] = (threadIdx.x * 1315423911ull + i) % q; Y[i] = (threadIdx.x * 2654435761ull + 7 * i) % q; }
    constexpr int SC = (MODE >= 20) ? 4 : 1;   // split chains (4 registers per residue)
    Split SX[SC], SY[SC];
#pragma unroll
    for (int i = 0; i < SC; ++i) { SX[i] = split_make(X[i]); SY[i] = split_make(Y[i]); }
    __syncthreads();
    u64 g0; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g0));
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < CH; ++i) {
            if (MODE >= 20) {   // FP64-assisted on the split representation (no moves), 21: csub every 2nd, 22: integer part only
                if (i >= SC) continue;
                const u64 t = dfmul_split<MODE == 22>(SY[i], w, ws, r0, r1, nq);
                u32 xl = (u32)__double2loint(SX[i].l), xh = (u32)__double2loint(SX[i].h);
                if (MODE == 21 && (i & 1)) {
                    const u64 x = csub(((u64)xh << 32) | xl, q4 + q4); xl = (u32)x; xh = (u32)(x >> 32);
                }
                split_add3(SX[i], xl, xh, t, q2, false);
                split_add3(SY[i], xl, xh, t, q4, true);
            } else
            if (MODE == 0) {          // current: exact shoup, exact csub, [0,4q)
                const u64 x = csub(X[i], q2), t = shoup_mul(Y[i], w, ws, nq);
                X[i] = x + t; Y[i] = x - t + q2;
            } else if (MODE == 1) {   // approx shoup, exact csub, [0,8q)
                const u64 x = csub(X[i], q4), t = shoup_a(Y[i], w, ws, nq);
                X[i] = x + t; Y[i] = x - t + q4;
            } else if (MODE == 2) {   // approx shoup, high-word csub
                const u64 x = csub_hi(X[i], q4), t = shoup_a(Y[i], w, ws, nq);
                X[i] = x + t; Y[i] = x - t + q4;
            } else if (MODE == 3) {   // approx shoup, no csub (timing floor only; values wrap)
                const u64 x = X[i], t = shoup_a(Y[i], w, ws, nq);
                X[i] = x + t; Y[i] = x - t + q4;
            } else if (MODE == 4) {   // inverse, current
                const u64 u = X[i], v = Y[i];
                X[i] = csub(u + v, q2); Y[i] = shoup_mul(u - v + q2, w, ws, nq);
            } else if (MODE == 5) {   // inverse, approx shoup + exact csub
                const u64 u = X[i], v = Y[i];
                X[i] = csub(u + v, q4); Y[i] = shoup_a(u - v + q4, w, ws, nq);
            } else if (MODE == 6) {   // inverse, approx shoup + high-word csub
                const u64 u = X[i], v = Y[i];
                X[i] = csub_hi(u + v, q4); Y[i] = shoup_a(u - v + q4, w, ws, nq);
            } else if (MODE == 9) {   // FP64-assisted, no csub
                const u64 t = dfmul(Y[i], w, ws, r0, r1, nq);
                const u64 x = X[i];
                X[i] = x + t + q2; Y[i] = x - t + q4;
            } else if (MODE == 10) {  // FP64-assisted, exact csub on every second butterfly (one per two stages)
                const u64 t = dfmul(Y[i], w, ws, r0, r1, nq);
                const u64 x = (i & 1) ? csub(X[i], q4 + q4) : X[i];
                X[i] = x + t + q2; Y[i] = x - t + q4;
            } else if (MODE == 11) {  // FP64-assisted, exact csub every butterfly
                const u64 t = dfmul(Y[i], w, ws, r0, r1, nq);
                const u64 x = csub(X[i], q4);
                X[i] = x + t + q2; Y[i] = x - t + q4;
            } else if (MODE == 12) {  // approx shoup, csub on every second butterfly
                const u64 x = (i & 1) ? csub(X[i], q4 + q4) : X[i], t = shoup_a(Y[i], w, ws, nq);
                X[i] = x + t; Y[i] = x - t + q4;
            } else if (MODE == 13) {  // FP64-assisted with I2F conversions, no csub
                const u64 t = dfmul_cvt(Y[i], w, ws, r0, r1, nq);
                const u64 x = X[i];
                X[i] = x + t + q2; Y[i] = x - t + q4;
            } else if (MODE == 14) {  // FP64-assisted with I2F conversions, csub every 2nd
                const u64 t = dfmul_cvt(Y[i], w, ws, r0, r1, nq);
                const u64 x = (i & 1) ? csub(X[i], q4 + q4) : X[i];
                X[i] = x + t + q2; Y[i] = x - t + q4;
            } else if (MODE == 15) {  // Montgomery-friendly prime (q = 1 mod 2^32), csub every 2nd
                const u64 x = (i & 1) ? csub(X[i], q4 + q4) : X[i], t = shoup_mf(Y[i], w, ws, (u32)(nq >> 32) + 1u);
                X[i] = x + t; Y[i] = x - t + q4;
            } else if (MODE == 16) {  // SURVEY 8d family (high word of -q = 0xF0000000), csub every 2nd
                const u64 x = (i & 1) ? csub(X[i], q4 + q4) : X[i], t = shoup_sol(Y[i], w, ws, nq);
                X[i] = x + t; Y[i] = x - t + q4;
            } else if (MODE == 17) {  // pointwise product as shipped: lazy single-mulhi Barrett, no final subtraction
                Y[i] = barrett_lazy(Y[i] & 0x1fffffffffffffffull, X[i] & 0x1fffffffffffffffull, nq, ws, 59) + (u64)i;
            } else if (MODE == 18) {  // pointwise product in Montgomery form
                Y[i] = mont_mul(Y[i] & 0x1fffffffffffffffull, X[i] & 0x1fffffffffffffffull, q, ws | 1) + (u64)i;
            } else if (MODE == 7) {   // shoup_a alone
                Y[i] = shoup_a(Y[i] + X[i], w, ws, nq);
            } else {                  // barrett product (pointwise)
                Y[i] = barrett_mul(csub(Y[i], q), csub(X[i], q), q, ws, 59);
            }
        }
    }
    long long t1 = clock64();
    u64 g1; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g1));
    u64 acc = 0;
#pragma unroll
    for (int i = 0; i < CH; ++i) acc += X[i] ^ Y[i] ^ split_join(SX[i % SC]) ^ split_join(SY[i % SC]);
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0) { cyc[blockIdx.x] = t1 - t0; cyc[gridDim.x + blockIdx.x] = (long long)(g1 - g0); }
}


// ---- FP64 butterfly for moduli below 2^50 (error-free product: h = y*w rounded, l = fma(y,w,-h) exact) ----------
template <int RED>
__global__ void __launch_bounds__(512, 2) kf(double* out, double q, double w, double wq) {
    double X[CH], Y[CH];
    const double M = 6755399441055744.0;   // 1.5 * 2^52: fma(y, wq, M) - M = rint(y * wq)
    const double qi = 1.0 / q;
#pragma unroll
    for (int i = 0; i < CH; ++i) { X[i] = (double)((threadIdx.x * 1315423911ull + i) % (1ull << 49)); Y[i] = (double)((threadIdx.x * 2654435761ull + 7 * i) % (1ull << 49)); }
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < CH; ++i) {
            const double y = Y[i];
            const double c = fma(y, wq, M) - M;
            const double h = y * w;
            const double l = fma(y, w, -h);
            const double d = fma(-c, q, h);
            const double T = d + l;
            double x = X[i];
            if (RED && (i & 3) == (it & 3)) x = fma(-(fma(x, qi, M) - M), q, x);   // one full reduction per 4 butterflies
            X[i] = x + T; Y[i] = x - T;
        }
    }
    double acc = 0;
#pragma unroll
    for (int i = 0; i < CH; ++i) acc += X[i] + Y[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}
template <int RED>
static void runf(const char* name, int sms, int threads = 512) {
    double* out;
    const int blocks = sms * 2 * 8;
    cudaMalloc(&out, sizeof(double) * blocks * 512);
    const double q = 1125899902124033.0, w = 93790016731446.0;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    kf<RED><<<blocks, threads>>>(out, q, w, w / q);
    cudaEventRecord(e0);
    kf<RED><<<blocks, threads>>>(out, q, w, w / q);
    cudaEventRecord(e1);
    cudaDeviceSynchronize();
    float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
    const double nb = (double)blocks * threads * ITERS * CH;
    const double clk_ev = (ms * 1e-3 * 1.965e9) / (nb / 32 / (sms * 4.0));
    printf("%-42s %6.2f Tbfly/s   %6.2f clk/warp-bfly/SMSP (events @1.965GHz)  err=%s\n", name, nb / (ms * 1e-3) * 1e-12, clk_ev,
           cudaGetErrorString(cudaGetLastError()));
    cudaFree(out);
}

template <int MODE>
static void run(const char* name, int sms, int threads = 512) {
    const u64 q = 1152921504606584833ull, w = 987813353222176621ull;
    const u64 ws = (u64)(((unsigned __int128)w << 64) / q);
    u64* out; long long* cyc;
    const int blocks = sms * 2 * 8;   // 8 full waves of 2 co-resident 512-thread CTAs per SM
    cudaMalloc(&out, sizeof(u64) * blocks * 512); cudaMalloc(&cyc, sizeof(long long) * blocks * 2);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    kb<MODE><<<blocks, threads>>>(out, cyc, q, w, ws);
    cudaEventRecord(e0);
    kb<MODE><<<blocks, threads>>>(out, cyc, q, w, ws);
    cudaEventRecord(e1);
    cudaDeviceSynchronize();
    float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
    long long* h = new long long[blocks * 2];
    cudaMemcpy(h, cyc, sizeof(long long) * blocks * 2, cudaMemcpyDeviceToHost);
    double mean = 0, ns = 0; for (int i = 0; i < blocks; ++i) { mean += h[i]; ns += h[blocks + i]; } mean /= blocks; ns /= blocks;
    const double mhz = mean / ns * 1e3;
    const double nb = (double)blocks * threads * ITERS * (MODE >= 20 ? 4 : CH);          // thread-butterflies
    const double per_s = nb / (ms * 1e-3);
    // whole-kernel figure: warp-butterflies per SMSP = nb/32/(sms*4); clocks at 1.965 GHz
    const double clk_ev = (ms * 1e-3 * 1.965e9) / (nb / 32 / (sms * 4.0));
    printf("%-42s %6.2f Tbfly/s  %6.2f clk/warp-bfly/SMSP (events @1.965GHz)  %6.2f (clock64, 8 warps/SMSP)  SM clock %4.0f MHz (clock64/globaltimer)  err=%s\n", name,
           per_s * 1e-12, clk_ev, mean / (8.0 * ITERS * CH), mhz, cudaGetErrorString(cudaGetLastError()));
    delete[] h; cudaFree(out); cudaFree(cyc);
}

int main(int argc, char** argv) {
    int sms = 0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    if (argc > 1) {   // single mode, for ncu
        const int m = atoi(argv[1]);
        if (m == 12) run<12>("approx shoup csub/2", sms);
        if (m == 20) run<20>("split no csub", sms);
        if (m == 21) run<21>("split csub/2", sms);
        if (m == 22) run<22>("split, integer part only", sms);
        if (m == 9) run<9>("fp64 no csub", sms);
        if (m == 3) run<3>("approx shoup no csub", sms);
        return 0;
    }
    run<0>("fwd  exact shoup + exact csub (current)", sms);
    run<0>("  same, 4 warps/SMSP (2 x 256 thr/SM)", sms, 256);
    run<0>("  same, 2 warps/SMSP (2 x 128 thr/SM)", sms, 128);
    run<12>("fwd approx shoup csub/2, 4 warps/SMSP", sms, 256);
    run<1>("fwd  approx shoup + exact csub", sms);
    run<2>("fwd  approx shoup + high-word csub", sms);
    run<3>("fwd  approx shoup, no csub (floor)", sms);
    run<4>("inv  exact shoup + exact csub (current)", sms);
    run<5>("inv  approx shoup + exact csub", sms);
    run<6>("inv  approx shoup + high-word csub", sms);
    run<9>("fwd  FP64-assisted, no csub", sms);
    run<10>("fwd  FP64-assisted, csub every 2nd", sms);
    run<11>("fwd  FP64-assisted, csub every bfly", sms);
    run<12>("fwd  approx shoup, csub every 2nd", sms);
    run<13>("fwd  FP64-assisted (I2F), no csub", sms);
    run<14>("fwd  FP64-assisted (I2F), csub every 2nd", sms);
    run<20>("fwd  FP64 split-representation, no csub", sms);
    run<21>("fwd  FP64 split-representation, csub/2", sms);
    run<20>("  same no csub, 4 warps/SMSP", sms, 256);
    run<22>("fwd  split, integer part only (fake quotient)", sms);
    runf<0>("FP64 butterfly q<2^50, no range reduction", sms);
    runf<1>("FP64 butterfly q<2^50, reduce 1 in 4", sms);
    runf<1>("  same, 4 warps/SMSP", sms, 256);
    run<15>("fwd  Montgomery-friendly q=1 mod 2^32, csub/2", sms);
    run<15>("  same, 4 warps/SMSP", sms, 256);
    run<16>("fwd  2^60-c family (shift for h0*nh), csub/2", sms);
    run<17>("pointwise: lazy Barrett (shipped)", sms);
    run<18>("pointwise: Montgomery form", sms);
    run<7>("shoup_a alone (+1 add)", sms);
    run<8>("barrett_mul (+2 csub)", sms);
    return 0;
}
