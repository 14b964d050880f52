// extern "C" boundary (include/gpufhe_b200.h): argument checking, error mapping, and the
// host-buffer entry points that pipeline H2D -> kernels -> D2H over limb groups.
#include <cstdio>
#include <cstdlib>
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstring>
#include <thread>
#include <vector>
#include "../../include/gpufhe_b200.h"
#include "pipe_sched.h"
#include "plan.h"

using namespace gfhe;

struct gfhe_plan { Plan impl; };   // never instantiated: gfhe_plan_t is a reinterpret of Plan*

static inline Plan* P(gfhe_plan_t h) { return reinterpret_cast<Plan*>(h); }
static inline cudaStream_t S(void* s) { return reinterpret_cast<cudaStream_t>(s); }

static int check_range(const Plan* pl, uint32_t limb0, uint32_t nl) {
    if (!pl) return GFHE_EARG;
    if (nl == 0 || (uint64_t)limb0 + nl > pl->L) return GFHE_ERANGE;
    return 0;
}

// The transform / product kernels move residues with 16-byte vector accesses and TMA bulk copies: a buffer that is not
// 16-byte aligned would raise a misaligned-address fault, which is a sticky context error.  Refuse it here instead.
static inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

extern "C" {

int gfhe_abi_version(void) { return GFHE_ABI_VERSION; }

const char* gfhe_error_string(int code) {
    switch (code) {
        case 0: return "success";
        case GFHE_EARG: return "null pointer or zero size";
        case GFHE_ELOGN: return "log2(N) outside [4, 20] or roots missing";
        case GFHE_EMODULUS: return "modulus outside [2, 2^62)";
        case GFHE_ENOTNTT: return "modulus is not 1 mod 2N";
        case GFHE_EROOT: return "psi is not a primitive 2N-th root of unity (psi^N != -1)";
        case GFHE_ENOMEM: return "host allocation failed";
        case GFHE_ERANGE: return "limb range outside the plan";
        case GFHE_EWORKSPACE: return "workspace missing or too small";
        case GFHE_ENOTABLES: return "plan has no NTT tables (created with log2_n = 0)";
        case GFHE_ETERMS: return "too many terms: terms * q^2 must fit 128 bits";
        case GFHE_EGALOIS: return "Galois element must be odd";
        case GFHE_ELAYOUT: return "packed 32-bit layout needs every modulus of the plan below 2^30";
        case GFHE_EALIGN: return "device buffer is not 16-byte aligned (transforms and products use 16-byte accesses)";
        case GFHE_EFORMAT: return "not a gpufhe_b200 plan image (bad magic, version, size or checksum)";
        case GFHE_ESTATICWS: return "gfhe_poly_mul_static: the job needs more scratch than the library's static workspace (GFHE_STATIC_WS_BYTES = 32 MiB)";
        case GFHE_ENOALLOC: return "gfhe_poly_mul_static: this plan runs the pipelined kernels (GFHE_PIPE=1), which allocate per stream";
        default: return code > 0 ? cudaGetErrorString((cudaError_t)code) : "unknown gpufhe_b200 error";
    }
}

int gfhe_plan_create(gfhe_plan_t* out, int device, uint32_t log2_n, uint32_t num_limbs,
                     const uint64_t* moduli, const uint64_t* psi) {
    if (!out) return GFHE_EARG;
    Plan* pl = nullptr;
    int rc = plan_create(&pl, device, log2_n, num_limbs, (const u64*)moduli, (const u64*)psi);
    if (rc == 0) *out = reinterpret_cast<gfhe_plan_t>(pl);
    return rc;
}

int gfhe_plan_destroy(gfhe_plan_t plan) { return plan_destroy(P(plan)); }

size_t gfhe_plan_export_bytes(gfhe_plan_t plan) { return plan ? plan_image_bytes(*P(plan)) : 0; }
int gfhe_plan_export(gfhe_plan_t plan, void* buf, size_t bytes) {
    if (!plan || !buf) return GFHE_EARG;
    return plan_export(*P(plan), buf, bytes);
}
int gfhe_plan_import(gfhe_plan_t* out, int device, const void* buf, size_t bytes) {
    if (!out || !buf) return GFHE_EARG;
    Plan* pl = nullptr;
    int rc = plan_import(&pl, device, buf, bytes);
    if (rc == 0) *out = reinterpret_cast<gfhe_plan_t>(pl);
    return rc;
}
int gfhe_plan_moduli(gfhe_plan_t plan, uint64_t* moduli, uint64_t* psi, uint32_t n) {
    if (!plan) return GFHE_EARG;
    const Plan* pl = P(plan);
    uint32_t k = 0;
    for (; k < n && k < pl->L; ++k) {
        if (moduli) moduli[k] = pl->moduli[k];
        if (psi) psi[k] = pl->logN ? pl->psi[k] : 0;
    }
    return (int)k;
}

int gfhe_plan_info(gfhe_plan_t plan, int* device, uint32_t* log2_n, uint32_t* num_limbs) {
    if (!plan) return GFHE_EARG;
    if (device) *device = P(plan)->device;
    if (log2_n) *log2_n = P(plan)->logN;
    if (num_limbs) *num_limbs = P(plan)->L;
    return 0;
}

int gfhe_debug_pipe_schedule(uint32_t passes, const uint32_t* tiles_per_pass, uint32_t npoly, uint32_t resident,
                             uint32_t* out, uint32_t max_tickets, uint32_t* lag) {
    if (!tiles_per_pass || passes < 2 || passes > 3 || npoly == 0) return GFHE_EARG;
    const PipeSched sc = make_pipe_sched(passes, tiles_per_pass, npoly, resident);
    if (lag) *lag = sc.lag;
    if (out)
        for (uint32_t t = 0; t < sc.total && t < max_tickets; ++t) pipe_decode(sc, t, out + 3 * t, out + 3 * t + 1, out + 3 * t + 2);
    return (int)sc.total;
}

int gfhe_profile_enable(gfhe_plan_t plan, int on) {
    if (!plan) return GFHE_EARG;
    Plan* pl = P(plan);
    if (!pl->prof) pl->prof = new Profiler();
    pl->prof->on = on != 0;
    return 0;
}

static const char* const kKernelNames[K_COUNT] = {"k_col_fwd", "k_row_fwd", "k_row_inv", "k_col_inv", "k_row_polymul", "k_vec", "k_mul_acc", "k_automorphism",
                                                  "k_pipe_ntt_fwd", "k_pipe_ntt_inv", "k_pipe_polymul"};

int gfhe_profile_read(gfhe_plan_t plan, uint32_t max_kernels, const char** names, double* total_ms, uint64_t* launches) {
    if (!plan) return GFHE_EARG;
    Plan* pl = P(plan);
    if (!pl->prof) return 0;
    DeviceGuard g(pl->device);
    double ms[K_COUNT] = {};
    uint64_t cnt[K_COUNT] = {};
    for (auto& r : pl->prof->recs) {
        cudaError_t e = cudaEventSynchronize(r.b);
        if (e != cudaSuccess) return -(int)e;
        float t = 0.f;
        if ((e = cudaEventElapsedTime(&t, r.a, r.b)) != cudaSuccess) return -(int)e;
        ms[r.kid] += t; cnt[r.kid] += 1;
        pl->prof->pool.push_back(r.a); pl->prof->pool.push_back(r.b);
    }
    pl->prof->recs.clear();
    uint32_t n = 0;
    for (int k = 0; k < K_COUNT && n < max_kernels; ++k) {
        if (!cnt[k]) continue;
        if (names) names[n] = kKernelNames[k];
        if (total_ms) total_ms[n] = ms[k];
        if (launches) launches[n] = cnt[k];
        ++n;
    }
    return (int)n;
}

int gfhe_ntt_forward(gfhe_plan_t plan, uint32_t limb0, uint32_t nl, uint32_t batch,
                     const uint64_t* d_in, uint64_t* d_out, void* stream) {
    if (int rc = check_range(P(plan), limb0, nl)) return rc;
    if (!d_in || !d_out || batch == 0) return GFHE_EARG;
    if (!aligned16(d_in) || !aligned16(d_out)) return GFHE_EALIGN;
    if (P(plan)->logN == 0) return GFHE_ENOTABLES;
    DeviceGuard g(P(plan)->device);
    return (int)launch_ntt_fwd(*P(plan), limb0, nl, (const u64*)d_in, (u64*)d_out, batch, S(stream));
}

int gfhe_ntt_inverse(gfhe_plan_t plan, uint32_t limb0, uint32_t nl, uint32_t batch,
                     const uint64_t* d_in, uint64_t* d_out, void* stream) {
    if (int rc = check_range(P(plan), limb0, nl)) return rc;
    if (!d_in || !d_out || batch == 0) return GFHE_EARG;
    if (!aligned16(d_in) || !aligned16(d_out)) return GFHE_EALIGN;
    if (P(plan)->logN == 0) return GFHE_ENOTABLES;
    DeviceGuard g(P(plan)->device);
    return (int)launch_ntt_inv(*P(plan), limb0, nl, (const u64*)d_in, (u64*)d_out, batch, S(stream));
}

size_t gfhe_poly_mul_workspace_bytes(gfhe_plan_t plan, uint32_t nl, uint32_t batch) {
    if (!plan || P(plan)->logN == 0) return 0;
    return poly_mul_workspace_bytes(*P(plan), nl, batch);
}

int gfhe_poly_mul(gfhe_plan_t plan, uint32_t limb0, uint32_t nl, uint32_t batch,
                  const uint64_t* d_a, const uint64_t* d_b, uint64_t* d_c,
                  void* d_ws, size_t ws_bytes, void* stream) {
    if (int rc = check_range(P(plan), limb0, nl)) return rc;
    if (!d_a || !d_b || !d_c || batch == 0) return GFHE_EARG;
    if (P(plan)->logN == 0) return GFHE_ENOTABLES;
    if (d_c == d_a || d_c == d_b) return GFHE_EARG;
    if (!aligned16(d_a) || !aligned16(d_b) || !aligned16(d_c) || !aligned16(d_ws)) return GFHE_EALIGN;
    const size_t need = poly_mul_workspace_bytes(*P(plan), nl, batch);
    if (need && (!d_ws || ws_bytes < need)) return GFHE_EWORKSPACE;
    DeviceGuard g(P(plan)->device);
    return (int)launch_poly_mul(*P(plan), limb0, nl, (const u64*)d_a, (const u64*)d_b, (u64*)d_c, (u64*)d_ws, batch,
                                S(stream));
}

static_assert(GFHE_STATIC_WS_BYTES == kStaticWsBytes, "header and library disagree on the static workspace");
size_t gfhe_static_workspace_bytes(void) { return kStaticWsBytes; }

// gfhe_poly_mul on the library's own scratch (plan.cu, g_static_ws): no cudaMalloc / cudaFree / stream or event creation
int gfhe_poly_mul_static(gfhe_plan_t plan, uint32_t limb0, uint32_t nl, uint32_t batch,
                         const uint64_t* d_a, const uint64_t* d_b, uint64_t* d_c, void* stream) {
    if (int rc = check_range(P(plan), limb0, nl)) return rc;
    if (!d_a || !d_b || !d_c || batch == 0) return GFHE_EARG;
    if (P(plan)->logN == 0) return GFHE_ENOTABLES;
    if (d_c == d_a || d_c == d_b) return GFHE_EARG;
    if (!aligned16(d_a) || !aligned16(d_b) || !aligned16(d_c)) return GFHE_EALIGN;
    if (P(plan)->pipe == 1) return GFHE_ENOALLOC;
    // overflow-safe: nl * batch * N * 8 against the array
    const unsigned long long words = (unsigned long long)nl * batch;
    if (P(plan)->logN > 12 && words > ((unsigned long long)kStaticWsBytes / sizeof(u64)) >> P(plan)->logN) return GFHE_ESTATICWS;
    DeviceGuard g(P(plan)->device);
    return (int)launch_poly_mul_static(*P(plan), limb0, nl, (const u64*)d_a, (const u64*)d_b, (u64*)d_c, batch, S(stream));
}

int gfhe_poly_mul_pre(gfhe_plan_t plan, uint32_t limb0, uint32_t nl, uint32_t batch,
                      const uint64_t* d_a, const uint64_t* d_b_ntt, uint64_t* d_c, void* stream) {
    if (int rc = check_range(P(plan), limb0, nl)) return rc;
    if (!d_a || !d_b_ntt || !d_c || batch == 0) return GFHE_EARG;
    if (P(plan)->logN == 0) return GFHE_ENOTABLES;
    if (d_c == d_a || d_c == d_b_ntt) return GFHE_EARG;
    if (!aligned16(d_a) || !aligned16(d_b_ntt) || !aligned16(d_c)) return GFHE_EALIGN;
    DeviceGuard g(P(plan)->device);
    return (int)launch_poly_mul_pre(*P(plan), limb0, nl, (const u64*)d_a, (const u64*)d_b_ntt, (u64*)d_c, batch, S(stream));
}

// ---- packed 32-bit layout (every modulus of the plan below 2^30) ----------------------------------------------------
static int check_packed(const Plan* pl, uint32_t limb0, uint32_t nl) {
    if (int rc = check_range(pl, limb0, nl)) return rc;
    if (pl->logN == 0) return GFHE_ENOTABLES;
    return pl->w32 ? 0 : GFHE_ELAYOUT;
}

int gfhe_ntt_forward_u32(gfhe_plan_t plan, uint32_t limb0, uint32_t nl, uint32_t batch,
                         const uint32_t* d_in, uint32_t* d_out, void* stream) {
    if (int rc = check_packed(P(plan), limb0, nl)) return rc;
    if (!d_in || !d_out || batch == 0) return GFHE_EARG;
    if (!aligned16(d_in) || !aligned16(d_out)) return GFHE_EALIGN;
    DeviceGuard g(P(plan)->device);
    return (int)launch_ntt_fwd_u32(*P(plan), limb0, nl, d_in, d_out, batch, S(stream));
}

int gfhe_ntt_inverse_u32(gfhe_plan_t plan, uint32_t limb0, uint32_t nl, uint32_t batch,
                         const uint32_t* d_in, uint32_t* d_out, void* stream) {
    if (int rc = check_packed(P(plan), limb0, nl)) return rc;
    if (!d_in || !d_out || batch == 0) return GFHE_EARG;
    if (!aligned16(d_in) || !aligned16(d_out)) return GFHE_EALIGN;
    DeviceGuard g(P(plan)->device);
    return (int)launch_ntt_inv_u32(*P(plan), limb0, nl, d_in, d_out, batch, S(stream));
}

int gfhe_poly_mul_u32(gfhe_plan_t plan, uint32_t limb0, uint32_t nl, uint32_t batch,
                      const uint32_t* d_a, const uint32_t* d_b, uint32_t* d_c,
                      void* d_ws, size_t ws_bytes, void* stream) {
    if (int rc = check_packed(P(plan), limb0, nl)) return rc;
    if (!d_a || !d_b || !d_c || batch == 0) return GFHE_EARG;
    if (d_c == d_a || d_c == d_b) return GFHE_EARG;
    if (!aligned16(d_a) || !aligned16(d_b) || !aligned16(d_c) || !aligned16(d_ws)) return GFHE_EALIGN;
    const size_t need = poly_mul_workspace_bytes(*P(plan), nl, batch) / 2;   // one operand-sized scratch of 32-bit words
    if (need && (!d_ws || ws_bytes < need)) return GFHE_EWORKSPACE;
    DeviceGuard g(P(plan)->device);
    return (int)launch_poly_mul_u32(*P(plan), limb0, nl, d_a, d_b, d_c, (uint32_t*)d_ws, batch, S(stream));
}

int gfhe_poly_mul_pre_u32(gfhe_plan_t plan, uint32_t limb0, uint32_t nl, uint32_t batch,
                          const uint32_t* d_a, const uint32_t* d_b_ntt, uint32_t* d_c, void* stream) {
    if (int rc = check_packed(P(plan), limb0, nl)) return rc;
    if (!d_a || !d_b_ntt || !d_c || batch == 0) return GFHE_EARG;
    if (d_c == d_a || d_c == d_b_ntt) return GFHE_EARG;
    if (!aligned16(d_a) || !aligned16(d_b_ntt) || !aligned16(d_c)) return GFHE_EALIGN;
    DeviceGuard g(P(plan)->device);
    return (int)launch_poly_mul_pre_u32(*P(plan), limb0, nl, d_a, d_b_ntt, d_c, batch, S(stream));
}

static int vec_op_i32(int op, gfhe_plan_t plan, uint32_t limb0, uint32_t nl, uint64_t inner,
                      const int32_t* a, const int32_t* b, int32_t* out, void* stream) {
    if (int rc = check_range(P(plan), limb0, nl)) return rc;
    if (!P(plan)->w32) return GFHE_ELAYOUT;
    if (inner == 0) return 0;
    if (!a || !b || !out) return GFHE_EARG;
    DeviceGuard g(P(plan)->device);
    return (int)launch_vec_op_u32(op, *P(plan), limb0, nl, a, b, out, inner, S(stream));
}
int gfhe_vec_add_i32(gfhe_plan_t p, uint32_t l0, uint32_t nl, uint64_t inner, const int32_t* a, const int32_t* b, int32_t* o, void* s) { return vec_op_i32(0, p, l0, nl, inner, a, b, o, s); }
int gfhe_vec_sub_i32(gfhe_plan_t p, uint32_t l0, uint32_t nl, uint64_t inner, const int32_t* a, const int32_t* b, int32_t* o, void* s) { return vec_op_i32(1, p, l0, nl, inner, a, b, o, s); }
int gfhe_vec_mul_i32(gfhe_plan_t p, uint32_t l0, uint32_t nl, uint64_t inner, const int32_t* a, const int32_t* b, int32_t* o, void* s) { return vec_op_i32(2, p, l0, nl, inner, a, b, o, s); }

int gfhe_ntt_mul_acc_u32(gfhe_plan_t plan, uint32_t limb0, uint32_t nl, uint64_t inner, uint32_t terms,
                         const uint32_t* d_a, const uint32_t* d_b, uint64_t term_stride, uint32_t* d_out, void* stream) {
    if (int rc = check_range(P(plan), limb0, nl)) return rc;
    if (!P(plan)->w32) return GFHE_ELAYOUT;
    if (inner == 0) return 0;
    if (!d_a || !d_b || !d_out || terms == 0) return GFHE_EARG;
    if (terms > 1 && (d_out == d_a || d_out == d_b)) return GFHE_EARG;
    if (terms > 65536u) return GFHE_ETERMS;
    DeviceGuard g(P(plan)->device);
    return (int)launch_mul_acc_u32(*P(plan), limb0, nl, inner, terms, d_a, d_b, term_stride, d_out, S(stream));
}

int gfhe_automorphism_u32(gfhe_plan_t plan, uint32_t limb0, uint32_t nl, uint32_t batch, uint32_t k, int ntt_domain,
                          const uint32_t* d_in, uint32_t* d_out, void* stream) {
    if (int rc = check_packed(P(plan), limb0, nl)) return rc;
    if (!d_in || !d_out || batch == 0 || d_in == d_out) return GFHE_EARG;
    if ((k & 1u) == 0) return GFHE_EGALOIS;
    DeviceGuard g(P(plan)->device);
    return (int)launch_automorphism_u32(*P(plan), limb0, nl, batch, k, ntt_domain != 0, d_in, d_out, S(stream));
}

int gfhe_ntt_mul_acc(gfhe_plan_t plan, uint32_t limb0, uint32_t nl, uint64_t inner, uint32_t terms,
                     const uint64_t* d_a, const uint64_t* d_b, uint64_t term_stride, uint64_t* d_out, void* stream) {
    if (int rc = check_range(P(plan), limb0, nl)) return rc;
    if (inner == 0) return 0;
    if (!d_a || !d_b || !d_out || terms == 0) return GFHE_EARG;
    if (terms > 1 && (d_out == d_a || d_out == d_b)) return GFHE_EARG;
    const uint32_t spare = 128u - 2u * P(plan)->max_bits;   // q < 2^62, so spare >= 4
    const uint32_t max_terms = spare >= 16 ? 65536u : (1u << spare);
    if (terms > max_terms) return GFHE_ETERMS;
    DeviceGuard g(P(plan)->device);
    return (int)launch_mul_acc(*P(plan), limb0, nl, inner, terms, (const u64*)d_a, (const u64*)d_b, term_stride, (u64*)d_out,
                               S(stream));
}

int gfhe_automorphism(gfhe_plan_t plan, uint32_t limb0, uint32_t nl, uint32_t batch, uint32_t k, int ntt_domain,
                      const uint64_t* d_in, uint64_t* d_out, void* stream) {
    if (int rc = check_range(P(plan), limb0, nl)) return rc;
    if (!d_in || !d_out || batch == 0 || d_in == d_out) return GFHE_EARG;
    if (P(plan)->logN == 0) return GFHE_ENOTABLES;
    if ((k & 1u) == 0) return GFHE_EGALOIS;
    DeviceGuard g(P(plan)->device);
    return (int)launch_automorphism(*P(plan), limb0, nl, batch, k, ntt_domain != 0, (const u64*)d_in, (u64*)d_out, S(stream));
}

static int vec_op(int op, gfhe_plan_t plan, uint32_t limb0, uint32_t nl, uint64_t inner,
                  const int64_t* a, const int64_t* b, int64_t* out, void* stream) {
    if (int rc = check_range(P(plan), limb0, nl)) return rc;
    if (inner == 0) return 0;
    if (!a || !b || !out) return GFHE_EARG;
    DeviceGuard g(P(plan)->device);
    return (int)launch_vec_op(op, *P(plan), limb0, nl, (const long long*)a, (const long long*)b, (long long*)out, inner,
                              S(stream));
}
int gfhe_vec_add(gfhe_plan_t p, uint32_t l0, uint32_t nl, uint64_t inner, const int64_t* a, const int64_t* b, int64_t* o, void* s) { return vec_op(0, p, l0, nl, inner, a, b, o, s); }
int gfhe_vec_sub(gfhe_plan_t p, uint32_t l0, uint32_t nl, uint64_t inner, const int64_t* a, const int64_t* b, int64_t* o, void* s) { return vec_op(1, p, l0, nl, inner, a, b, o, s); }
int gfhe_vec_mul(gfhe_plan_t p, uint32_t l0, uint32_t nl, uint64_t inner, const int64_t* a, const int64_t* b, int64_t* o, void* s) { return vec_op(2, p, l0, nl, inner, a, b, o, s); }
int gfhe_vec_add_u64(gfhe_plan_t p, uint32_t l0, uint32_t nl, uint64_t inner, const uint64_t* a, const uint64_t* b, uint64_t* o, void* s) { return vec_op(4, p, l0, nl, inner, (const int64_t*)a, (const int64_t*)b, (int64_t*)o, s); }
int gfhe_vec_sub_u64(gfhe_plan_t p, uint32_t l0, uint32_t nl, uint64_t inner, const uint64_t* a, const uint64_t* b, uint64_t* o, void* s) { return vec_op(5, p, l0, nl, inner, (const int64_t*)a, (const int64_t*)b, (int64_t*)o, s); }
int gfhe_vec_mul_u64(gfhe_plan_t p, uint32_t l0, uint32_t nl, uint64_t inner, const uint64_t* a, const uint64_t* b, uint64_t* o, void* s) { return vec_op(6, p, l0, nl, inner, (const int64_t*)a, (const int64_t*)b, (int64_t*)o, s); }

}  // extern "C"

// ------------------------------------------------------------------------------------------------
// host-buffer pipeline: limb groups flow H2D (stream 0) -> kernels (stream 1) -> D2H (stream 2)
// through three device slots, so PCIe in, compute and PCIe out overlap.
//
// Pinned host arrays (cudaHostAlloc / torch pin_memory) are copied straight from / to the caller's
// pointers.  Pageable arrays -- what a numpy caller of the reference actually holds -- would make
// cudaMemcpyAsync synchronous and serialise the three streams, so they are staged through an
// internal ring of pinned buffers (one per slot and operand): a few host threads copy group g+1 into
// its staging buffers while group g crosses PCIe, and drain finished results the same way.
// ------------------------------------------------------------------------------------------------
namespace gfhe {

struct HostPipe {
    static constexpr int kSlots = 3;
    cudaStream_t s_in = nullptr, s_run = nullptr, s_out = nullptr;
    cudaEvent_t in_done[kSlots] = {}, run_done[kSlots] = {}, out_done[kSlots] = {};
    u64* buf[kSlots][4] = {};   // device: a, b, c, workspace
    size_t cap = 0;             // bytes per device buffer
    u64* pin[kSlots][3] = {};   // pinned staging (pageable callers only): a, b, c
    size_t pin_cap = 0;
    std::vector<cudaEvent_t> tev;   // timing events, six per group: H2D / kernels / D2H begin and end
    double stats[GFHE_HOST_STATS_N] = {};
};

void host_pipe_destroy(Plan* p) {
    HostPipe* hp = static_cast<HostPipe*>(p->host_pipe);
    if (!hp) return;
    for (int s = 0; s < HostPipe::kSlots; ++s) {
        for (int k = 0; k < 4; ++k) if (hp->buf[s][k]) cudaFree(hp->buf[s][k]);
        for (int k = 0; k < 3; ++k) if (hp->pin[s][k]) cudaFreeHost(hp->pin[s][k]);
        if (hp->in_done[s]) cudaEventDestroy(hp->in_done[s]);
        if (hp->run_done[s]) cudaEventDestroy(hp->run_done[s]);
        if (hp->out_done[s]) cudaEventDestroy(hp->out_done[s]);
    }
    for (cudaEvent_t e : hp->tev) cudaEventDestroy(e);
    if (hp->s_in) cudaStreamDestroy(hp->s_in);
    if (hp->s_run) cudaStreamDestroy(hp->s_run);
    if (hp->s_out) cudaStreamDestroy(hp->s_out);
    delete hp;
    p->host_pipe = nullptr;
}

static cudaError_t host_pipe_get(Plan* p, size_t bytes, bool need_ws, bool staged, size_t ngroups, HostPipe** out) {
    HostPipe* hp = static_cast<HostPipe*>(p->host_pipe);
    cudaError_t e = cudaSuccess;
    if (!hp) {
        hp = new HostPipe();
        p->host_pipe = hp;
        if ((e = cudaStreamCreateWithFlags(&hp->s_in, cudaStreamNonBlocking))) return e;
        if ((e = cudaStreamCreateWithFlags(&hp->s_run, cudaStreamNonBlocking))) return e;
        if ((e = cudaStreamCreateWithFlags(&hp->s_out, cudaStreamNonBlocking))) return e;
        for (int s = 0; s < HostPipe::kSlots; ++s) {
            if ((e = cudaEventCreateWithFlags(&hp->in_done[s], cudaEventDisableTiming))) return e;
            if ((e = cudaEventCreateWithFlags(&hp->run_done[s], cudaEventDisableTiming))) return e;
            if ((e = cudaEventCreateWithFlags(&hp->out_done[s], cudaEventDisableTiming))) return e;
        }
    }
    if (hp->cap < bytes) {
        for (int s = 0; s < HostPipe::kSlots; ++s)
            for (int k = 0; k < 4; ++k) {
                if (hp->buf[s][k]) cudaFree(hp->buf[s][k]);
                hp->buf[s][k] = nullptr;
                if ((e = cudaMalloc(&hp->buf[s][k], bytes))) { hp->cap = 0; return e; }
            }
        hp->cap = bytes;
    }
    (void)need_ws;
    if (staged && hp->pin_cap < bytes) {
        for (int s = 0; s < HostPipe::kSlots; ++s)
            for (int k = 0; k < 3; ++k) {
                if (hp->pin[s][k]) cudaFreeHost(hp->pin[s][k]);
                hp->pin[s][k] = nullptr;
                if ((e = cudaHostAlloc(&hp->pin[s][k], bytes, cudaHostAllocDefault))) { hp->pin_cap = 0; return e; }
            }
        hp->pin_cap = bytes;
    }
    while (hp->tev.size() < 6 * ngroups) {
        cudaEvent_t ev;
        if ((e = cudaEventCreate(&ev))) return e;
        hp->tev.push_back(ev);
    }
    *out = hp;
    return cudaSuccess;
}

// true when the driver can DMA straight from / to p (pinned or managed); false for plain pageable memory
static bool host_ptr_is_pinned(const void* p) {
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return at.type == cudaMemoryTypeHost || at.type == cudaMemoryTypeManaged;
}

// memcpy split over a few host threads (one core moves ~10 GB/s; PCIe Gen5 x16 wants ~50)
static void par_memcpy(void* dst, const void* src, size_t bytes, unsigned nthr) {
    if (nthr <= 1 || bytes < (4u << 20)) { std::memcpy(dst, src, bytes); return; }
    std::vector<std::thread> pool;
    const size_t chunk = ((bytes / nthr) + 4095) & ~(size_t)4095;
    for (unsigned t = 1; t < nthr; ++t) {
        const size_t o = (size_t)t * chunk;
        if (o >= bytes) break;
        const size_t n = bytes - o < chunk ? bytes - o : chunk;
        pool.emplace_back([=]() { std::memcpy((char*)dst + o, (const char*)src + o, n); });
    }
    std::memcpy(dst, src, chunk < bytes ? chunk : bytes);
    for (auto& th : pool) th.join();
}

static inline double now_ms() {
    return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// op: 0 forward NTT, 1 inverse NTT, 2 poly-mul
static int run_host(Plan* pl, int op, u32 batch, const u64* h_a, const u64* h_b, u64* h_c) {
    if (!pl || !h_a || !h_c || batch == 0 || (op == 2 && !h_b)) return GFHE_EARG;
    if (pl->logN == 0) return GFHE_ENOTABLES;
    DeviceGuard g(pl->device);
    const double t_begin = now_ms();
    const size_t limb_bytes = ((size_t)batch * sizeof(u64)) << pl->logN;
    // limb groups of ~32 MiB per operand: large enough for full PCIe rate, small enough to overlap
    size_t group_mib = 32;
    // a job of a few large limbs (at most 32 MiB per operand: the config-4 shard of one of 8 GPUs is 4 limbs of 8 MiB) would be
    // 3 groups; 4 MiB groups cut its batch into 8 (measured on 8 ranks: 5.22 -> 5.03 ms per call, profiles/r02_run26_groups_n8.txt)
    if (limb_bytes * pl->L <= ((size_t)32 << 20)) group_mib = 4;
    if (const char* env = std::getenv("GFHE_HOST_GROUP_MIB")) { const long v = std::atol(env); if (v >= 1 && v <= 4096) group_mib = (size_t)v; }
    u32 gl = (u32)(((group_mib << 20) + limb_bytes - 1) / limb_bytes);
    if (gl < 1) gl = 1;
    if (gl > pl->L) gl = pl->L;
    // pageable operands are staged through pinned buffers (GFHE_HOST_STAGING=0 hands them to the driver as they are)
    bool staged = !(host_ptr_is_pinned(h_a) && host_ptr_is_pinned(h_c) && (op != 2 || host_ptr_is_pinned(h_b)));
    if (const char* env = std::getenv("GFHE_HOST_STAGING")) if (env[0] == '0') staged = false;
    unsigned nthr = std::thread::hardware_concurrency() / 2;   // staging threads: half the cores, 2..8
    if (nthr < 2) nthr = 2;
    if (nthr > 8) nthr = 8;
    if (const char* env = std::getenv("GFHE_HOST_COPY_THREADS")) { const long v = std::atol(env); if (v >= 1 && v <= 64) nthr = (unsigned)v; }
    const size_t limb_elems = (size_t)batch << pl->logN;
    // group schedule: full groups, then a halving tail (gl/2, gl/4, ..., 1) so that what is left after the last H2D
    // copy -- the last group's kernels and its D2H copy -- is one limb's worth, not a full group's
    // A group size below one limb's bytes cuts every limb's batch instead (a small shard -- 4 limbs on one of 8 GPUs -- is
    // otherwise 3 groups, and the pipeline's fill and drain are most of the call).
    struct Group { u32 l0, nl, b0, nb; };
    std::vector<Group> groups;
    const u32 parts = (u32)std::min<size_t>(batch, (limb_bytes + (group_mib << 20) - 1) / (group_mib << 20));
    if (parts >= 2) {
        for (u32 l = 0; l < pl->L; ++l)
            for (u32 k = 0; k < parts; ++k) {
                const u32 b0 = (u32)((u64)batch * k / parts), b1 = (u32)((u64)batch * (k + 1) / parts);
                groups.push_back({l, 1u, b0, b1 - b0});
            }
    } else {
        u32 left = pl->L, l0 = 0;
        while (left > gl) { groups.push_back({l0, gl, 0u, batch}); l0 += gl; left -= gl; }
        for (u32 gsz = left; left > 0;) { gsz = gsz > 1 ? (gsz + 1) / 2 : 1; if (gsz > left) gsz = left; groups.push_back({l0, gsz, 0u, batch}); l0 += gsz; left -= gsz; }
    }
    const size_t poly_elems = (size_t)1 << pl->logN, poly_bytes = sizeof(u64) << pl->logN;
    auto g_off = [&](const Group& g) { return limb_elems * g.l0 + poly_elems * g.b0; };          // words from the start of a host operand
    auto g_bytes = [&](const Group& g) { return poly_bytes * g.nb * g.nl; };
    HostPipe* hp = nullptr;
    cudaError_t e = host_pipe_get(pl, limb_bytes * gl, op == 2, staged, groups.size(), &hp);
    if (e) return (int)e;
    constexpr int kS = HostPipe::kSlots;
    double stage_in_ms = 0, stage_out_ms = 0;
    // staged mode: a drain thread copies finished groups out of their pinned slots (in order) while this thread stages
    // the next inputs in; `issued` / `drained` are the hand-shake (a slot is reused only after its group was drained)
    std::atomic<size_t> issued{0}, drained{0};
    std::atomic<int> drain_err{0};
    std::atomic<bool> abort_drain{false};
    const unsigned in_thr = nthr, out_thr = nthr > 2 ? nthr / 2 : 1;
    std::thread drainer;
    if (staged)
        drainer = std::thread([&]() {
            cudaSetDevice(pl->device);
            for (size_t g2 = 0; g2 < groups.size(); ++g2) {
                while (issued.load(std::memory_order_acquire) <= g2) {
                    if (abort_drain.load(std::memory_order_acquire)) return;
                    std::this_thread::yield();
                }
                const int slot = (int)(g2 % kS);
                if (cudaError_t de = cudaEventSynchronize(hp->out_done[slot])) { drain_err.store((int)de); return; }
                const double t0 = now_ms();
                par_memcpy(h_c + g_off(groups[g2]), hp->pin[slot][2], g_bytes(groups[g2]), out_thr);
                stage_out_ms += now_ms() - t0;
                drained.store(g2 + 1, std::memory_order_release);
            }
        });
    auto body = [&]() -> cudaError_t {
        for (size_t it = 0; it < groups.size(); ++it) {
            const int slot = (int)(it % kS);
            const u32 nl = groups[it].nl, l0 = groups[it].l0, nb = groups[it].nb;
            const size_t bytes = g_bytes(groups[it]), off = g_off(groups[it]);
            u64 *da = hp->buf[slot][0], *db = hp->buf[slot][1], *dc = hp->buf[slot][2], *dw = hp->buf[slot][3];
            const u64 *src_a = h_a + off, *src_b = h_b ? h_b + off : nullptr;
            u64* dst_c = h_c + off;
            if (staged) {
                if (it >= (size_t)kS)   // the slot's three staging buffers are free once its previous group was drained
                    while (drained.load(std::memory_order_acquire) < it - kS + 1) {
                        if (drain_err.load()) return (cudaError_t)drain_err.load();
                        std::this_thread::yield();
                    }
                const double t0 = now_ms();
                par_memcpy(hp->pin[slot][0], src_a, bytes, in_thr);
                if (op == 2) par_memcpy(hp->pin[slot][1], src_b, bytes, in_thr);
                stage_in_ms += now_ms() - t0;
                src_a = hp->pin[slot][0]; src_b = hp->pin[slot][1]; dst_c = hp->pin[slot][2];
            }
            if (it >= (size_t)kS) {   // slot reuse: inputs may be overwritten once its kernels ran, output once copied out
                if ((e = cudaStreamWaitEvent(hp->s_in, hp->run_done[slot], 0))) return e;
                if ((e = cudaStreamWaitEvent(hp->s_run, hp->out_done[slot], 0))) return e;
            }
            cudaEvent_t* tv = hp->tev.data() + 6 * it;
            if ((e = cudaEventRecord(tv[0], hp->s_in))) return e;
            if ((e = cudaMemcpyAsync(da, src_a, bytes, cudaMemcpyHostToDevice, hp->s_in))) return e;
            if (op == 2 && (e = cudaMemcpyAsync(db, src_b, bytes, cudaMemcpyHostToDevice, hp->s_in))) return e;
            if ((e = cudaEventRecord(tv[1], hp->s_in))) return e;
            if ((e = cudaEventRecord(hp->in_done[slot], hp->s_in))) return e;
            if ((e = cudaStreamWaitEvent(hp->s_run, hp->in_done[slot], 0))) return e;
            if ((e = cudaEventRecord(tv[2], hp->s_run))) return e;
            if (op == 0) e = launch_ntt_fwd(*pl, l0, nl, da, dc, nb, hp->s_run);
            else if (op == 1) e = launch_ntt_inv(*pl, l0, nl, da, dc, nb, hp->s_run);
            else e = launch_poly_mul(*pl, l0, nl, da, db, dc, dw, nb, hp->s_run);
            if (e) return e;
            if ((e = cudaEventRecord(tv[3], hp->s_run))) return e;
            if ((e = cudaEventRecord(hp->run_done[slot], hp->s_run))) return e;
            if ((e = cudaStreamWaitEvent(hp->s_out, hp->run_done[slot], 0))) return e;
            if ((e = cudaEventRecord(tv[4], hp->s_out))) return e;
            if ((e = cudaMemcpyAsync(dst_c, dc, bytes, cudaMemcpyDeviceToHost, hp->s_out))) return e;
            if ((e = cudaEventRecord(tv[5], hp->s_out))) return e;
            if ((e = cudaEventRecord(hp->out_done[slot], hp->s_out))) return e;
            issued.store(it + 1, std::memory_order_release);
        }
        return cudaSuccess;
    };
    e = body();
    if (staged) {
        if (e != cudaSuccess) abort_drain.store(true, std::memory_order_release);   // groups that were never issued will not arrive
        drainer.join();
        if (e == cudaSuccess && drain_err.load()) e = (cudaError_t)drain_err.load();
    }
    // success or not, nothing of this call may still be in flight on the caller's buffers when it returns
    const cudaError_t e1 = cudaStreamSynchronize(hp->s_out), e2 = cudaStreamSynchronize(hp->s_run), e3 = cudaStreamSynchronize(hp->s_in);
    if (e == cudaSuccess) e = e1 ? e1 : (e2 ? e2 : e3);
    if (e) return (int)e;
    double busy[3] = {0, 0, 0};
    for (size_t it = 0; it < issued.load(); ++it)
        for (int k = 0; k < 3; ++k) {
            float ms = 0.f;
            if (cudaEventElapsedTime(&ms, hp->tev[6 * it + 2 * k], hp->tev[6 * it + 2 * k + 1]) == cudaSuccess) busy[k] += ms;
        }
    const double total = (double)limb_bytes * pl->L;
    hp->stats[GFHE_HOST_STAT_H2D_MS] = busy[0];
    hp->stats[GFHE_HOST_STAT_KERNEL_MS] = busy[1];
    hp->stats[GFHE_HOST_STAT_D2H_MS] = busy[2];
    hp->stats[GFHE_HOST_STAT_H2D_BYTES] = total * (op == 2 ? 2 : 1);
    hp->stats[GFHE_HOST_STAT_D2H_BYTES] = total;
    hp->stats[GFHE_HOST_STAT_WALL_MS] = now_ms() - t_begin;
    hp->stats[GFHE_HOST_STAT_STAGE_IN_MS] = stage_in_ms;
    hp->stats[GFHE_HOST_STAT_STAGE_OUT_MS] = stage_out_ms;
    hp->stats[GFHE_HOST_STAT_STAGED] = staged ? 1.0 : 0.0;
    hp->stats[GFHE_HOST_STAT_GROUPS] = (double)groups.size();
    return 0;
}

}  // namespace gfhe

extern "C" {
int gfhe_ntt_host(gfhe_plan_t plan, int inverse, uint32_t batch, const uint64_t* h_in, uint64_t* h_out) {
    return run_host(P(plan), inverse ? 1 : 0, batch, (const u64*)h_in, nullptr, (u64*)h_out);
}
int gfhe_poly_mul_host(gfhe_plan_t plan, uint32_t batch, const uint64_t* h_a, const uint64_t* h_b, uint64_t* h_c) {
    return run_host(P(plan), 2, batch, (const u64*)h_a, (const u64*)h_b, (u64*)h_c);
}
int gfhe_host_stats(gfhe_plan_t plan, double* out, uint32_t n) {
    if (!plan || !out) return GFHE_EARG;
    const HostPipe* hp = static_cast<const HostPipe*>(P(plan)->host_pipe);
    uint32_t k = 0;
    for (; k < n && k < GFHE_HOST_STATS_N; ++k) out[k] = hp ? hp->stats[k] : 0.0;
    return (int)k;
}
}
