// Which form of the 32x32 multiply is cheapest on sm_100a, and which of them block the ALU pipe?
// Follow-up to pipebench.cu (IMAD.WIDE + IADD3 do not overlap there).  Event-timed, 8 waves.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/mulbench tools/mulbench.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
typedef unsigned long long u64; typedef unsigned int u32;
constexpr int ITERS = 4096, CH = 8;

template <int OP>
__global__ void __launch_bounds__(512, 2) k(u64* out, u32 seed) {
    u32 a[CH], b[CH], c[CH]; u64 w[CH]; double d[CH];
#pragma unroll
    for (int i = 0; i < CH; ++i) { a[i] = seed + threadIdx.x * 7 + i; b[i] = seed * 3 + i; c[i] = seed ^ i; w[i] = ((u64)a[i] << 32) | b[i]; d[i] = (double)a[i]; }
    const u32 m = seed | 1, n = seed ^ 0x9e3779b9u;
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < CH; ++i) {
            if (OP == 0) asm volatile("{.reg .u32 l,h; mov.b64 {l,h}, %0; mul.wide.u32 %0, l, h;}" : "+l"(w[i]));                       // W, no addend
            if (OP == 1) asm volatile("{.reg .u32 l,h; mov.b64 {l,h}, %0; mad.wide.u32 %0, l, %1, %0;}" : "+l"(w[i]) : "r"(m));           // W, 64-bit addend
            if (OP == 2) { asm volatile("{.reg .u32 l,h; mov.b64 {l,h}, %0; mul.wide.u32 %0, l, h;}" : "+l"(w[i]));
                           asm volatile("{.reg .u32 t; add.u32 t, %0, %1; add.u32 %0, t, %2;}" : "+r"(c[i]) : "r"(m), "r"(n)); }             // W + IADD3
            if (OP == 3) { asm volatile("{.reg .u32 l,h; mov.b64 {l,h}, %0; mad.wide.u32 %0, l, %1, %0;}" : "+l"(w[i]) : "r"(m));
                           asm volatile("{.reg .u32 t; add.u32 t, %0, %1; add.u32 %0, t, %2;}" : "+r"(c[i]) : "r"(m), "r"(n)); }             // WA + IADD3
            if (OP == 4) asm volatile("mul.hi.u32 %0, %0, %1;" : "+r"(a[i]) : "r"(m));                                                        // H
            if (OP == 5) asm volatile("mad.hi.u32 %0, %0, %1, %0;" : "+r"(a[i]) : "r"(m));                                                    // HA
            if (OP == 6) asm volatile("mad.lo.u32 %0, %0, %1, %0;" : "+r"(a[i]) : "r"(m));                                                    // L
            if (OP == 7) { asm volatile("{.reg .u32 l,h; mov.b64 {l,h}, %0; mul.wide.u32 %0, l, h;}" : "+l"(w[i]));
                           asm volatile("mad.lo.u32 %0, %0, %1, %0;" : "+r"(a[i]) : "r"(m)); }                                                // W + L
            if (OP == 8) { asm volatile("{.reg .u32 l,h; mov.b64 {l,h}, %0; mul.wide.u32 %0, l, h;}" : "+l"(w[i]));
                           asm volatile("{.reg .u32 t; add.u32 t, %0, %1; add.u32 %0, t, %2;}" : "+r"(c[i]) : "r"(m), "r"(n));
                           asm volatile("{.reg .u32 t; add.u32 t, %0, %1; add.u32 %0, t, %2;}" : "+r"(b[i]) : "r"(m), "r"(n)); }             // W + 2 IADD3
            if (OP == 9) { asm volatile("mad.lo.u32 %0, %0, %1, %0;" : "+r"(a[i]) : "r"(m));
                           asm volatile("{.reg .u32 t; add.u32 t, %0, %1; add.u32 %0, t, %2;}" : "+r"(c[i]) : "r"(m), "r"(n)); }             // L + IADD3
            if (OP == 10) { asm volatile("mul.hi.u32 %0, %0, %1;" : "+r"(a[i]) : "r"(m));
                            asm volatile("{.reg .u32 t; add.u32 t, %0, %1; add.u32 %0, t, %2;}" : "+r"(c[i]) : "r"(m), "r"(n)); }            // H + IADD3
            if (OP == 11) { asm volatile("{.reg .u32 l,h; mov.b64 {l,h}, %0; mul.wide.u32 %0, l, h;}" : "+l"(w[i]));
                            asm volatile("mad.lo.u32 %0, %0, %1, %0;" : "+r"(a[i]) : "r"(m));
                            asm volatile("{.reg .u32 t; add.u32 t, %0, %1; add.u32 %0, t, %2;}" : "+r"(c[i]) : "r"(m), "r"(n));
                            asm volatile("{.reg .u32 t; add.u32 t, %0, %1; add.u32 %0, t, %2;}" : "+r"(b[i]) : "r"(m), "r"(n)); }            // W + L + 2 IADD3
            if (OP == 12) asm volatile("{.reg .u32 t; add.u32 t, %0, %1; add.u32 %0, t, %2;}" : "+r"(c[i]) : "r"(m), "r"(n));               // IADD3
            if (OP == 13) { asm volatile("{.reg .u32 l,h; mov.b64 {l,h}, %0; mul.wide.u32 %0, l, h;}" : "+l"(w[i]));
                            asm volatile("fma.rn.f64 %0, %0, %1, %0;" : "+d"(d[i]) : "d"(1.0000001)); }                                       // W + DFMA
            if (OP == 14) asm volatile("fma.rn.f64 %0, %0, %1, %0;" : "+d"(d[i]) : "d"(1.0000001));                                          // DFMA
            if (OP == 15) { asm volatile("mad.lo.u32 %0, %0, %1, %0;" : "+r"(a[i]) : "r"(m));
                            asm volatile("fma.rn.f64 %0, %0, %1, %0;" : "+d"(d[i]) : "d"(1.0000001)); }                                       // L + DFMA
        }
    }
    u64 acc = 0;
#pragma unroll
    for (int i = 0; i < CH; ++i) acc += a[i] + b[i] + c[i] + w[i] + (u64)d[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

template <int OP>
static void run(const char* name, int sms, int nops) {
    u64* out;
    const int blocks = sms * 2 * 4;
    cudaMalloc(&out, sizeof(u64) * blocks * 512);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<OP><<<blocks, 512>>>(out, 12345u);
    cudaEventRecord(e0);
    k<OP><<<blocks, 512>>>(out, 12345u);
    cudaEventRecord(e1);
    cudaDeviceSynchronize();
    float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
    const double groups = (double)blocks * 16 * ITERS * CH / (sms * 4.0);
    const double clk = ms * 1e-3 * 1965e6 / groups;
    printf("%-44s %6.2f clk per group per SMSP (%d instr, %5.2f each)  err=%s\n", name, clk, nops, clk / nops, cudaGetErrorString(cudaGetLastError()));
    cudaFree(out);
}

int main() {
    int sms = 0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    run<0>("W  (IMAD.WIDE, no addend)", sms, 1);
    run<1>("WA (IMAD.WIDE, 64-bit addend)", sms, 1);
    run<6>("L  (IMAD lo)", sms, 1);
    run<4>("H  (IMAD.HI)", sms, 1);
    run<5>("HA (IMAD.HI with addend)", sms, 1);
    run<12>("IADD3", sms, 1);
    run<14>("DFMA", sms, 1);
    run<2>("W + IADD3", sms, 2);
    run<3>("WA + IADD3", sms, 2);
    run<8>("W + 2 IADD3", sms, 3);
    run<9>("L + IADD3", sms, 2);
    run<10>("H + IADD3", sms, 2);
    run<7>("W + L", sms, 2);
    run<11>("W + L + 2 IADD3", sms, 4);
    run<13>("W + DFMA", sms, 2);
    run<15>("L + DFMA", sms, 2);
    return 0;
}
