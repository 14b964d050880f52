// Pipe-rate microbenchmark for sm_100a (B200): warp-instruction issue cost per SM sub-partition of
// the integer / FP64 instructions a 64-bit modular butterfly is made of, alone and in pairs.
// Timed with CUDA events over 8 full waves (clock64 of one warp per CTA is NOT a fair sample:
// the warp arbiter is priority-based), converted to clocks at the SM clock read from NVML-free
// cudaDevAttrClockRate.   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipebench tools/pipebench.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
typedef unsigned long long u64; typedef unsigned int u32;

constexpr int ITERS = 4096, CH = 8;

template <int OP>
__global__ void __launch_bounds__(512, 2) k(u64* out, u32 seed) {
    u32 a[CH], b[CH]; u64 w[CH]; double d[CH]; float f[CH];
#pragma unroll
    for (int i = 0; i < CH; ++i) { a[i] = seed + threadIdx.x * 7 + i; b[i] = seed * 3 + i; w[i] = ((u64)a[i] << 32) | b[i]; d[i] = a[i]; f[i] = b[i]; }
    const u32 m = seed | 1, n = seed ^ 0x9e3779b9u;
    const u64 mn = (u64)m << 32 | n;
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < CH; ++i) {
            if (OP == 0 || OP == 7 || OP == 10) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(m), "r"(n));
            if (OP == 1 || OP == 11 || OP == 12) asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(w[i]) : "r"(b[i]), "r"(m));
            if (OP == 2) asm volatile("mad.hi.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(m), "r"(n));
            if (OP == 3 || OP == 7 || OP == 11) asm volatile("{.reg .u32 t; add.u32 t, %0, %1; add.u32 %0, t, %2;}" : "+r"(b[i]) : "r"(m), "r"(n));
            if (OP == 4) asm volatile("add.u64 %0, %0, %1;" : "+l"(w[i]) : "l"(mn));
            if (OP == 5 || OP == 10 || OP == 12) asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(d[i]) : "d"(1.0000001), "d"(0.5));
            if (OP == 6) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[i]) : "r"(m), "r"(n));
            if (OP == 8) asm volatile("shf.l.wrap.b32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(m), "r"(5));
            if (OP == 9) asm volatile("{.reg .pred p; setp.lt.u32 p, %0, %1; selp.u32 %0, %2, %0, p;}" : "+r"(a[i]) : "r"(m), "r"(n));
            if (OP == 13) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(f[i]) : "f"(1.0000001f), "f"(0.5f));
            if (OP == 14) asm volatile("cvt.rn.f64.u64 %0, %1;" : "=d"(d[i]) : "l"(w[i] + (u64)d[i]));
            if (OP == 15) asm volatile("cvt.rzi.u64.f64 %0, %1;" : "=l"(w[i]) : "d"(d[i] + (double)w[i]));
            if (OP == 20) asm volatile("cvt.rn.f64.u32 %0, %1;" : "=d"(d[i]) : "r"(a[i] + (u32)__double2loint(d[i])));
            if (OP == 16) asm volatile("mul.lo.u64 %0, %0, %1;" : "+l"(w[i]) : "l"(mn));
            if (OP == 17) asm volatile("mul.hi.u64 %0, %0, %1;" : "+l"(w[i]) : "l"(mn));
            if (OP == 18) asm volatile("add.f64 %0, %0, %1;" : "+d"(d[i]) : "d"(0.5));
            if (OP == 19) asm volatile("{.reg .u32 t; mul.hi.u32 t, %0, %1; mad.lo.u32 %0, %0, %1, t;}" : "+r"(a[i]) : "r"(m));
        }
    }
    u64 acc = 0;
#pragma unroll
    for (int i = 0; i < CH; ++i) acc += a[i] + b[i] + w[i] + (u64)d[i] + (u64)f[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

static double g_mhz = 1965.0;
template <int OP>
static void run(const char* name, int sms, int ops_per_iter) {
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
    const double warp_instr_per_smsp = (double)blocks * 16 * ITERS * CH * ops_per_iter / (sms * 4.0);
    const double clk = ms * 1e-3 * g_mhz * 1e6 / warp_instr_per_smsp;
    printf("%-34s %6.2f clk per warp-instr per SMSP   %6.1f lanes/clk/SM   err=%s\n", name, clk, 128.0 / clk,
           cudaGetErrorString(cudaGetLastError()));
    cudaFree(out);
}

int main(int argc, char** argv) {
    int sms = 0, khz = 0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    if (argc > 1) g_mhz = atof(argv[1]); else if (khz > 0) g_mhz = khz / 1000.0;
    printf("SMs %d, assumed SM clock %.0f MHz (pass the observed clock as argv[1] to override)\n", sms, g_mhz);
    run<0>("IMAD (mad.lo.u32)", sms, 1);
    run<1>("IMAD.WIDE.U32 (64-bit acc)", sms, 1);
    run<2>("IMAD.HI.U32", sms, 1);
    run<19>("mul.hi + mad.lo pair", sms, 2);
    run<3>("IADD3", sms, 1);
    run<4>("add.u64 (IADD3 + IADD3.X)", sms, 2);
    run<6>("LOP3", sms, 1);
    run<8>("SHF", sms, 1);
    run<9>("ISETP + SEL", sms, 2);
    run<13>("FFMA", sms, 1);
    run<5>("DFMA", sms, 1);
    run<18>("DADD", sms, 1);
    run<7>("IMAD + IADD3 interleaved", sms, 2);
    run<11>("IMAD.WIDE + IADD3 interleaved", sms, 2);
    run<10>("IMAD + DFMA interleaved", sms, 2);
    run<12>("IMAD.WIDE + DFMA interleaved", sms, 2);
    run<14>("I2F.F64.U64 (+add,+cvt back)", sms, 1);
    run<15>("F2I.U64.F64 (+add,+cvt back)", sms, 1);
    run<20>("I2F.F64.U32 (+1 IADD)", sms, 1);
    run<16>("mul.lo.u64 (as compiled)", sms, 1);
    run<17>("mul.hi.u64 (as compiled)", sms, 1);
    return 0;
}
