// Host-side plan: per-limb constants and twiddle tables, generated once on the host from the
// primitive 2N-th roots supplied by primitive.py and kept resident in HBM.
#pragma once
#include <cuda_runtime.h>
#include <mutex>
#include <unordered_map>
#include <vector>
#include "modarith.cuh"

namespace gfhe {

// Optional per-kernel timing (gfhe_profile_enable): every launch is bracketed by cudaEvents on
// its own stream, so bench.py can report each kernel's live duration and share of the step.
enum KernelId { K_COL_FWD = 0, K_ROW_FWD, K_ROW_INV, K_COL_INV, K_ROW_POLYMUL, K_VEC, K_MAC, K_AUTO, K_PIPE_FWD, K_PIPE_INV, K_PIPE_MUL, K_COUNT };
struct Profiler {
    bool on = false;
    struct Rec { int kid; cudaEvent_t a, b; };
    std::vector<Rec> recs;
    std::vector<cudaEvent_t> pool;
    cudaEvent_t get() {
        if (!pool.empty()) { cudaEvent_t e = pool.back(); pool.pop_back(); return e; }
        cudaEvent_t e; cudaEventCreate(&e); return e;
    }
};

struct Plan {
    int device = 0;
    u32 logN = 0;                      // 0: modulus set only (vec_* ops), no tables
    u32 L = 0;
    u32 max_bits = 0;                  // bit length of the largest modulus
    bool f64 = false;                  // every modulus is below 2^50 (and not all below 2^30): the FP64 kernels and FP64-form tables
    bool w32 = false;                  // every modulus is below 2^30: the Shoup32 kernels are used
    bool lazy = false;                 // every modulus is in [2^19, 2^60): the Lazy60 kernels are used
    int persist = -1;                  // fused row pass: -1 = per-policy default, 0 / 1 = GFHE_PERSIST at plan creation
    int pipe = -1;                     // N > 4096: one pipelined launch with the intermediate in L2 (ntt_pipe.inl); -1 = default (on), 0 / 1 = GFHE_PIPE
    LimbParams* d_params = nullptr;    // [L]
    ulonglong2* d_fwd = nullptr;       // [L][N] {psi^brv(k), Shoup companion}; entry 0 unused
    ulonglong2* d_inv = nullptr;       // [L][N] {psi^-brv(k), Shoup companion}
    ulonglong2* d_fwd_t0 = nullptr;    // [L][N/512][15][32] warp-tiled copy of d_fwd for the window at bit 0
    ulonglong2* d_inv_t0 = nullptr;    // [L][N/512][15][32] same for d_inv (ntt_core.cuh, TwT0)
    std::vector<u64> moduli, psi;      // host copies (psi empty when logN == 0): what gfhe_plan_moduli / the plan image report
    int split = -1;                    // two-stream split of small jobs (plan.cu, split_run): -1 = by size, 0 / 1 = GFHE_SPLIT
    mutable cudaStream_t aux_stream = nullptr;      // second stream of split_run, with its fork / join events
    mutable cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    void* host_pipe = nullptr;         // staging buffers/streams of the *_host entry points (capi.cu)
    // synchronisation words of the pipelined kernels, one buffer per stream that launched one (left zeroed by every launch)
    struct SyncBuf { u32* d = nullptr; size_t words = 0; };
    mutable std::mutex sync_mu;
    mutable std::unordered_map<cudaStream_t, SyncBuf> sync_bufs;
    Profiler* prof = nullptr;
};

struct ProfScope {
    Profiler* p; cudaStream_t st; size_t idx;
    ProfScope(const Plan& pl, int kid, cudaStream_t s) : p(pl.prof && pl.prof->on ? pl.prof : nullptr), st(s), idx(0) {
        if (!p || p->recs.size() >= 65536) { p = nullptr; return; }
        Profiler::Rec r{kid, p->get(), p->get()};
        cudaEventRecord(r.a, st);
        idx = p->recs.size();
        p->recs.push_back(r);
    }
    ~ProfScope() { if (p) cudaEventRecord(p->recs[idx].b, st); }
};

// The kernels (and the plan's allocations) live on the plan's device whatever the caller's current device is; the caller's
// device is restored on scope exit, so torch's notion of the current device never changes behind its back.
struct DeviceGuard {
    int prev = -1;
    explicit DeviceGuard(int dev) { cudaGetDevice(&prev); if (prev != dev) cudaSetDevice(dev); else prev = -1; }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

// plan.cu
int plan_create(Plan** out, int device, u32 logN, u32 L, const u64* moduli, const u64* psi);
int plan_destroy(Plan* p);
void plan_read_env(Plan* pl);   // GFHE_PERSIST / GFHE_PIPE / GFHE_SPLIT
// zeroed buffer of at least `words` 32-bit words for a pipelined launch on stream st (stream-ordered reuse)
cudaError_t pipe_sync_get(const Plan& pl, cudaStream_t st, size_t words, u32** out);
// Serialisable table cache: a plan image is the header + moduli + roots + the four device
// tables exactly as they sit in HBM, so importing it is validation + four uploads -- no root search, no table generation.
size_t plan_image_bytes(const Plan& pl);
int plan_export(const Plan& pl, void* buf, size_t bytes);
int plan_import(Plan** out, int device, const void* buf, size_t bytes);

// ntt_kernels.inl, instantiated per arithmetic policy by ntt_harvey.cu / ntt_lazy.cu
template <class A>
struct Launcher {
    typedef typename A::store S;   // word type of the arrays in HBM (u64; u32 for the packed 32-bit layout)
    static cudaError_t ntt_fwd(const Plan& pl, u32 limb0, u32 nl, const S* in, S* out, u32 B, cudaStream_t st);
    static cudaError_t ntt_inv(const Plan& pl, u32 limb0, u32 nl, const S* in, S* out, u32 B, cudaStream_t st);
    static cudaError_t poly_mul(const Plan& pl, u32 limb0, u32 nl, const S* a, const S* b, S* c, S* ws, u32 B, cudaStream_t st);
    static cudaError_t poly_mul_pre(const Plan& pl, u32 limb0, u32 nl, const S* a, const S* b_ntt, S* c, u32 B, cudaStream_t st);
};

// plan.cu -- pick the policy of the plan; all operate on limbs [limb0, limb0 + nl), buffers are [nl, B, N]
cudaError_t launch_ntt_fwd(const Plan& pl, u32 limb0, u32 nl, const u64* in, u64* out, u32 B, cudaStream_t st);
cudaError_t launch_ntt_inv(const Plan& pl, u32 limb0, u32 nl, const u64* in, u64* out, u32 B, cudaStream_t st);
cudaError_t launch_poly_mul(const Plan& pl, u32 limb0, u32 nl, const u64* a, const u64* b, u64* c, u64* ws, u32 B,
                            cudaStream_t st);
size_t poly_mul_workspace_bytes(const Plan& pl, u32 nl, u32 B);
cudaError_t launch_poly_mul_pre(const Plan& pl, u32 limb0, u32 nl, const u64* a, const u64* b_ntt, u64* c, u32 B, cudaStream_t st);
// the same product on the library's file-scope scratch array (plan.cu): allocates nothing, single stream
constexpr size_t kStaticWsBytes = (size_t)32 << 20;
cudaError_t launch_poly_mul_static(const Plan& pl, u32 limb0, u32 nl, const u64* a, const u64* b, u64* c, u32 B, cudaStream_t st);
// packed 32-bit layout (plans whose moduli are all below 2^30): the same operations on arrays of 32-bit words
cudaError_t launch_ntt_fwd_u32(const Plan& pl, u32 limb0, u32 nl, const u32* in, u32* out, u32 B, cudaStream_t st);
cudaError_t launch_ntt_inv_u32(const Plan& pl, u32 limb0, u32 nl, const u32* in, u32* out, u32 B, cudaStream_t st);
cudaError_t launch_poly_mul_u32(const Plan& pl, u32 limb0, u32 nl, const u32* a, const u32* b, u32* c, u32* ws, u32 B, cudaStream_t st);
cudaError_t launch_poly_mul_pre_u32(const Plan& pl, u32 limb0, u32 nl, const u32* a, const u32* b_ntt, u32* c, u32 B, cudaStream_t st);

// vec_kernels.cu
cudaError_t launch_vec_op(int op, const Plan& pl, u32 limb0, u32 nl, const long long* a, const long long* b,
                          long long* out, u64 inner, cudaStream_t st);

// packed_kernels.cu -- the operations around the transforms on the packed 32-bit layout (plans with every modulus below 2^30)
cudaError_t launch_vec_op_u32(int op, const Plan& pl, u32 limb0, u32 nl, const int* a, const int* b, int* out, u64 inner,
                              cudaStream_t st);
cudaError_t launch_mul_acc_u32(const Plan& pl, u32 limb0, u32 nl, u64 inner, u32 terms, const u32* a, const u32* b,
                               u64 term_stride, u32* out, cudaStream_t st);
cudaError_t launch_automorphism_u32(const Plan& pl, u32 limb0, u32 nl, u32 B, u32 k, bool ntt_domain, const u32* in, u32* out,
                                    cudaStream_t st);

// ring_kernels.cu -- NTT-domain multiply-accumulate and the Galois automorphism X -> X^k
cudaError_t launch_mul_acc(const Plan& pl, u32 limb0, u32 nl, u64 inner, u32 terms, const u64* a, const u64* b,
                           u64 term_stride, u64* out, cudaStream_t st);
cudaError_t launch_automorphism(const Plan& pl, u32 limb0, u32 nl, u32 B, u32 k, bool ntt_domain, const u64* in, u64* out,
                                cudaStream_t st);

}  // namespace gfhe
