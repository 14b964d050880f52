// extern "C" boundary (include/gpufhe_b200.h): argument checking, error mapping, and the
// host-buffer entry points that pipeline H2D -> kernels -> D2H over limb groups.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include "../../include/gpufhe_b200.h"
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

// The kernels are launched on the plan's device whatever the caller's current device is.
struct DeviceGuard {
    int prev = -1;
    explicit DeviceGuard(int dev) { cudaGetDevice(&prev); if (prev != dev) cudaSetDevice(dev); else prev = -1; }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

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

int gfhe_plan_info(gfhe_plan_t plan, int* device, uint32_t* log2_n, uint32_t* num_limbs) {
    if (!plan) return GFHE_EARG;
    if (device) *device = P(plan)->device;
    if (log2_n) *log2_n = P(plan)->logN;
    if (num_limbs) *num_limbs = P(plan)->L;
    return 0;
}

int gfhe_profile_enable(gfhe_plan_t plan, int on) {
    if (!plan) return GFHE_EARG;
    Plan* pl = P(plan);
    if (!pl->prof) pl->prof = new Profiler();
    pl->prof->on = on != 0;
    return 0;
}

static const char* const kKernelNames[K_COUNT] = {"k_col_fwd", "k_row_fwd", "k_row_inv", "k_col_inv", "k_row_polymul", "k_vec", "k_mul_acc", "k_automorphism"};

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
    if (P(plan)->logN == 0) return GFHE_ENOTABLES;
    DeviceGuard g(P(plan)->device);
    return (int)launch_ntt_fwd(*P(plan), limb0, nl, (const u64*)d_in, (u64*)d_out, batch, S(stream));
}

int gfhe_ntt_inverse(gfhe_plan_t plan, uint32_t limb0, uint32_t nl, uint32_t batch,
                     const uint64_t* d_in, uint64_t* d_out, void* stream) {
    if (int rc = check_range(P(plan), limb0, nl)) return rc;
    if (!d_in || !d_out || batch == 0) return GFHE_EARG;
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
    const size_t need = poly_mul_workspace_bytes(*P(plan), nl, batch);
    if (need && (!d_ws || ws_bytes < need)) return GFHE_EWORKSPACE;
    DeviceGuard g(P(plan)->device);
    return (int)launch_poly_mul(*P(plan), limb0, nl, (const u64*)d_a, (const u64*)d_b, (u64*)d_c, (u64*)d_ws, batch,
                                S(stream));
}

int gfhe_poly_mul_pre(gfhe_plan_t plan, uint32_t limb0, uint32_t nl, uint32_t batch,
                      const uint64_t* d_a, const uint64_t* d_b_ntt, uint64_t* d_c, void* stream) {
    if (int rc = check_range(P(plan), limb0, nl)) return rc;
    if (!d_a || !d_b_ntt || !d_c || batch == 0) return GFHE_EARG;
    if (P(plan)->logN == 0) return GFHE_ENOTABLES;
    if (d_c == d_a || d_c == d_b_ntt) return GFHE_EARG;
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
    DeviceGuard g(P(plan)->device);
    return (int)launch_ntt_fwd_u32(*P(plan), limb0, nl, d_in, d_out, batch, S(stream));
}

int gfhe_ntt_inverse_u32(gfhe_plan_t plan, uint32_t limb0, uint32_t nl, uint32_t batch,
                         const uint32_t* d_in, uint32_t* d_out, void* stream) {
    if (int rc = check_packed(P(plan), limb0, nl)) return rc;
    if (!d_in || !d_out || batch == 0) return GFHE_EARG;
    DeviceGuard g(P(plan)->device);
    return (int)launch_ntt_inv_u32(*P(plan), limb0, nl, d_in, d_out, batch, S(stream));
}

int gfhe_poly_mul_u32(gfhe_plan_t plan, uint32_t limb0, uint32_t nl, uint32_t batch,
                      const uint32_t* d_a, const uint32_t* d_b, uint32_t* d_c,
                      void* d_ws, size_t ws_bytes, void* stream) {
    if (int rc = check_packed(P(plan), limb0, nl)) return rc;
    if (!d_a || !d_b || !d_c || batch == 0) return GFHE_EARG;
    if (d_c == d_a || d_c == d_b) return GFHE_EARG;
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

}  // extern "C"

// ------------------------------------------------------------------------------------------------
// host-buffer pipeline: limb groups flow H2D (stream 0) -> kernels (stream 1) -> D2H (stream 2)
// through two device slots, so PCIe in, compute and PCIe out overlap.
// ------------------------------------------------------------------------------------------------
namespace gfhe {

struct HostPipe {
    static constexpr int kSlots = 2;
    cudaStream_t s_in = nullptr, s_run = nullptr, s_out = nullptr;
    cudaEvent_t in_done[kSlots] = {}, run_done[kSlots] = {}, out_done[kSlots] = {};
    u64* buf[kSlots][4] = {};   // a, b, c, workspace
    size_t cap = 0;             // bytes per buffer
};

void host_pipe_destroy(Plan* p) {
    HostPipe* hp = static_cast<HostPipe*>(p->host_pipe);
    if (!hp) return;
    for (int s = 0; s < HostPipe::kSlots; ++s) {
        for (int k = 0; k < 4; ++k) if (hp->buf[s][k]) cudaFree(hp->buf[s][k]);
        if (hp->in_done[s]) cudaEventDestroy(hp->in_done[s]);
        if (hp->run_done[s]) cudaEventDestroy(hp->run_done[s]);
        if (hp->out_done[s]) cudaEventDestroy(hp->out_done[s]);
    }
    if (hp->s_in) cudaStreamDestroy(hp->s_in);
    if (hp->s_run) cudaStreamDestroy(hp->s_run);
    if (hp->s_out) cudaStreamDestroy(hp->s_out);
    delete hp;
    p->host_pipe = nullptr;
}

static cudaError_t host_pipe_get(Plan* p, size_t bytes, HostPipe** out) {
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
    *out = hp;
    return cudaSuccess;
}

// op: 0 forward NTT, 1 inverse NTT, 2 poly-mul
static int run_host(Plan* pl, int op, u32 batch, const u64* h_a, const u64* h_b, u64* h_c) {
    if (!pl || !h_a || !h_c || batch == 0 || (op == 2 && !h_b)) return GFHE_EARG;
    if (pl->logN == 0) return GFHE_ENOTABLES;
    DeviceGuard g(pl->device);
    const size_t limb_bytes = ((size_t)batch * sizeof(u64)) << pl->logN;
    // limb groups of ~32 MiB per operand: large enough for full PCIe rate, small enough to overlap
    size_t group_mib = 32;
    if (const char* env = std::getenv("GFHE_HOST_GROUP_MIB")) { const long v = std::atol(env); if (v >= 1 && v <= 4096) group_mib = (size_t)v; }
    u32 gl = (u32)(((group_mib << 20) + limb_bytes - 1) / limb_bytes);
    if (gl < 1) gl = 1;
    if (gl > pl->L) gl = pl->L;
    HostPipe* hp = nullptr;
    cudaError_t e = host_pipe_get(pl, limb_bytes * gl, &hp);
    if (e) return (int)e;
    const size_t limb_elems = (size_t)batch << pl->logN;
    // group schedule: full groups, then a halving tail (gl/2, gl/4, ..., 1) so that what is left after the last H2D
    // copy -- the last group's kernels and its D2H copy -- is one limb's worth, not a full group's
    std::vector<u32> groups;
    {
        u32 left = pl->L;
        while (left > gl) { groups.push_back(gl); left -= gl; }
        for (u32 g = left; left > 0;) { g = g > 1 ? (g + 1) / 2 : 1; if (g > left) g = left; groups.push_back(g); left -= g; }
    }
    int slot = 0;
    u32 l0 = 0;
    for (size_t it = 0; it < groups.size(); l0 += groups[it], ++it, slot ^= 1) {
        const u32 nl = groups[it];
        const size_t bytes = limb_bytes * nl, off = limb_elems * l0;
        u64 *da = hp->buf[slot][0], *db = hp->buf[slot][1], *dc = hp->buf[slot][2], *dw = hp->buf[slot][3];
        if (it >= 2) {   // slot reuse: inputs may be overwritten once its kernels ran, output once copied out
            if ((e = cudaStreamWaitEvent(hp->s_in, hp->run_done[slot], 0))) return (int)e;
            if ((e = cudaStreamWaitEvent(hp->s_run, hp->out_done[slot], 0))) return (int)e;
        }
        if ((e = cudaMemcpyAsync(da, h_a + off, bytes, cudaMemcpyHostToDevice, hp->s_in))) return (int)e;
        if (op == 2 && (e = cudaMemcpyAsync(db, h_b + off, bytes, cudaMemcpyHostToDevice, hp->s_in))) return (int)e;
        if ((e = cudaEventRecord(hp->in_done[slot], hp->s_in))) return (int)e;
        if ((e = cudaStreamWaitEvent(hp->s_run, hp->in_done[slot], 0))) return (int)e;
        if (op == 0) e = launch_ntt_fwd(*pl, l0, nl, da, dc, batch, hp->s_run);
        else if (op == 1) e = launch_ntt_inv(*pl, l0, nl, da, dc, batch, hp->s_run);
        else e = launch_poly_mul(*pl, l0, nl, da, db, dc, dw, batch, hp->s_run);
        if (e) return (int)e;
        if ((e = cudaEventRecord(hp->run_done[slot], hp->s_run))) return (int)e;
        if ((e = cudaStreamWaitEvent(hp->s_out, hp->run_done[slot], 0))) return (int)e;
        if ((e = cudaMemcpyAsync(h_c + off, dc, bytes, cudaMemcpyDeviceToHost, hp->s_out))) return (int)e;
        if ((e = cudaEventRecord(hp->out_done[slot], hp->s_out))) return (int)e;
    }
    if ((e = cudaStreamSynchronize(hp->s_out))) return (int)e;
    if ((e = cudaStreamSynchronize(hp->s_run))) return (int)e;
    if ((e = cudaStreamSynchronize(hp->s_in))) return (int)e;
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
}
