// Integer-pipe microbenchmark for sm_100a: how many IMAD / IMAD.WIDE / IMAD.HI / IADD3 / DFMA
// warp-lanes retire per clock per SM, and what one Shoup/Harvey butterfly costs.  These rates
// bound the 64-bit NTT and are quoted in DESIGN.md.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipebench tools/pipebench.cu && ./pipebench
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../gpu-fhe_b200/csrc/modarith.cuh"
using namespace gfhe;

constexpr int ITERS = 2048, CH = 8;

template <int OP>
__global__ void __launch_bounds__(512) k(u64* out, long long* cyc, u32 seed) {
    u32 a[CH], b[CH]; u64 w[CH]; double d[CH];
#pragma unroll
    for (int i = 0; i < CH; ++i) { a[i] = seed + threadIdx.x * 7 + i; b[i] = seed * 3 + i; w[i] = ((u64)a[i] << 32) | b[i]; d[i] = a[i]; }
    const u32 m = seed | 1, n = seed ^ 0x9e3779b9u;
    __syncthreads();
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < CH; ++i) {
            if (OP == 0) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(m), "r"(n));
            if (OP == 1) asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(w[i]) : "r"(a[i]), "r"(m));
            if (OP == 2) asm volatile("mad.hi.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(m), "r"(n));
            if (OP == 3) asm volatile("{.reg .u32 t; add.u32 t, %0, %1; add.u32 %0, t, %2;}" : "+r"(a[i]) : "r"(m), "r"(n));   // one IADD3
            if (OP == 4) asm volatile("{.reg .u32 lo, hi, l2, h2; mov.b64 {lo,hi}, %0; mov.b64 {l2,h2}, %1; add.cc.u32 lo, lo, l2; addc.u32 hi, hi, h2; mov.b64 %0, {lo,hi};}" : "+l"(w[i]) : "l"((u64)m << 32 | n));
            if (OP == 5) asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(d[i]) : "d"(1.0000001), "d"(0.5));
            if (OP == 6) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[i]) : "r"(m), "r"(n));
            if (OP == 7) { asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(m), "r"(n)); asm volatile("{.reg .u32 t; add.u32 t, %0, %1; add.u32 %0, t, %2;}" : "+r"(b[i]) : "r"(m), "r"(n)); }
            if (OP == 8) asm volatile("shf.l.wrap.b32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(m), "r"(5));
            if (OP == 9) asm volatile("{.reg .pred p; setp.lt.u32 p, %0, %1; selp.u32 %0, %2, %0, p;}" : "+r"(a[i]) : "r"(m), "r"(n));
        }
    }
    long long t1 = clock64();
    u64 acc = 0;
#pragma unroll
    for (int i = 0; i < CH; ++i) acc += a[i] + b[i] + w[i] + (u64)d[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

// full butterflies on 8 independent (X,Y) pairs per thread
template <int MODE>
__global__ void __launch_bounds__(512) kb(u64* out, long long* cyc, u64 q, u64 w, u64 ws) {
    u64 X[CH], Y[CH];
    const u64 q2 = 2 * q, nq = 0 - q;
#pragma unroll
    for (int i = 0; i < CH; ++i) { X[i] = (threadIdx.x * 1315423911ull + i) % q; Y[i] = (threadIdx.x * 2654435761ull + 7 * i) % q; }
    __syncthreads();
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITERS / 4; ++it) {
#pragma unroll
        for (int i = 0; i < CH; ++i) {
            if (MODE == 0) {          // forward CT (Harvey): csub + shoup + add + sub
                const u64 x = csub(X[i], q2), t = shoup_mul(Y[i], w, ws, nq);
                X[i] = x + t; Y[i] = x - t + q2;
            } else if (MODE == 1) {   // inverse GS
                const u64 u = X[i], v = Y[i];
                X[i] = csub(u + v, q2); Y[i] = shoup_mul(u - v + q2, w, ws, nq);
            } else {                  // shoup_mul alone
                Y[i] = shoup_mul(Y[i] + X[i], w, ws, nq);
            }
        }
    }
    long long t1 = clock64();
    u64 acc = 0;
#pragma unroll
    for (int i = 0; i < CH; ++i) acc += X[i] ^ Y[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <typename F>
static void run(const char* name, F launch, int sms, double ops_per_thread, int ops_per_iter) {
    u64* out; long long* cyc;
    const int blocks = sms * 2;
    cudaMalloc(&out, sizeof(u64) * blocks * 512); cudaMalloc(&cyc, sizeof(long long) * blocks);
    launch(blocks, out, cyc); launch(blocks, out, cyc);
    cudaDeviceSynchronize();
    long long* h = new long long[blocks];
    cudaMemcpy(h, cyc, sizeof(long long) * blocks, cudaMemcpyDeviceToHost);
    double mean = 0; for (int i = 0; i < blocks; ++i) mean += h[i]; mean /= blocks;
    const double per_clk_sm = 1024.0 * ops_per_thread / mean;
    printf("%-28s %8.1f lanes/clk/SM   (%.2f issue-clk per warp-instr per SMSP)  err=%s\n", name, per_clk_sm * ops_per_iter,
           128.0 / (per_clk_sm * ops_per_iter), cudaGetErrorString(cudaGetLastError()));
    delete[] h; cudaFree(out); cudaFree(cyc);
}

int main() {
    int sms = 0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    printf("device %s, %d SMs, L2 %d MB\n", p.name, sms, p.l2CacheSize >> 20);
    const double n = (double)ITERS * CH;
#define RUN(OP, NAME, MULT) run(NAME, [](int b, u64* o, long long* c) { k<OP><<<b, 512>>>(o, c, 12345u); }, sms, n, MULT)
    RUN(0, "IMAD (mad.lo.u32)", 1);
    RUN(1, "IMAD.WIDE.U32", 1);
    RUN(2, "IMAD.HI.U32", 1);
    RUN(3, "IADD3", 1);
    RUN(4, "IADD3+IADD3.X (64b add)", 2);
    RUN(5, "DFMA", 1);
    RUN(6, "LOP3", 1);
    RUN(7, "IMAD + IADD3 interleaved", 2);
    RUN(8, "SHF", 1);
    RUN(9, "ISETP+SEL", 2);
    const u64 q = 1152921504606584833ull, w = 987813353222176621ull;
    const u64 ws = (u64)(((unsigned __int128)w << 64) / q);
    const double nb = (double)(ITERS / 4) * CH;
    run("butterfly fwd (per bfly)", [=](int b, u64* o, long long* c) { kb<0><<<b, 512>>>(o, c, q, w, ws); }, sms, nb, 1);
    run("butterfly inv (per bfly)", [=](int b, u64* o, long long* c) { kb<1><<<b, 512>>>(o, c, q, w, ws); }, sms, nb, 1);
    run("shoup_mul alone", [=](int b, u64* o, long long* c) { kb<2><<<b, 512>>>(o, c, q, w, ws); }, sms, nb, 1);
    return 0;
}
