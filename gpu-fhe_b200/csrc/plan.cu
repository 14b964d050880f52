// Plan construction: twiddle and Shoup-companion tables are generated ONCE on the host from the
// primitive 2N-th roots that primitive.py found, uploaded, and stay resident in HBM.
//
// The reference has no such code (/root/reference/primitive.py:3-4 is a placeholder); the
// table layout is dictated by the kernels in ntt_kernels.cu:
//   fwd[l][k] = { psi_l^brv(k),  floor(psi_l^brv(k)  * 2^64 / q_l) }   k in [1, N)
//   inv[l][k] = { psi_l^-brv(k), floor(psi_l^-brv(k) * 2^64 / q_l) }
// brv = bit reversal over log2(N) bits, so stage "m groups" reads entries [m, 2m) contiguously.
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>
#include "plan.h"

namespace gfhe {

typedef unsigned __int128 u128;

static inline u64 h_mulmod(u64 a, u64 b, u64 q) { return (u64)((u128)a * b % q); }
static u64 h_powmod(u64 b, u64 e, u64 q) {
    u64 r = 1 % q;
    b %= q;
    for (; e; e >>= 1) {
        if (e & 1) r = h_mulmod(r, b, q);
        b = h_mulmod(b, b, q);
    }
    return r;
}
static inline u64 h_shoup(u64 w, u64 q) { return (u64)(((u128)w << 64) / q); }
static inline u32 h_brv(u32 k, u32 bits) {
    u32 r = 0;
    for (u32 i = 0; i < bits; ++i) { r = (r << 1) | (k & 1); k >>= 1; }
    return r;
}

static void fill_limb_params(LimbParams& P, u64 q, u32 logN, u64 psi) {
    P = LimbParams{};
    P.q = q;
    P.q2 = 2 * q;
    P.nq = 0 - q;
    u32 k = 64 - (u32)__builtin_clzll(q);             // bit length
    P.bar_shift = k - 1;
    u128 mu = ((u128)1 << (k + 63)) / q;               // < 2^64 unless q is a power of two
    P.bar_mu = mu >> 64 ? ~0ull : (u64)mu;
    P.bar64 = (u64)((((u128)1) << 64) / q);
    P.r64 = (u64)((((u128)1) << 64) % q);
    if (k >= 20) {
        P.kshift = k - 17;
        P.kmu = (u64)((((u128)1) << (32 + P.kshift)) / q);
    }
    P.qd = (double)q;
    P.qinv = 1.0 / (double)q;
    if (logN) {
        const u64 N = 1ull << logN;
        P.ninv = h_powmod(N % q, q - 2, q);
        P.ninv_s = h_shoup(P.ninv, q);
        const u64 psi_inv = h_powmod(psi, 2 * N - 1, q);
        const u64 w1 = h_powmod(psi_inv, N >> 1, q);   // psi^-brv(1) = psi^-(N/2)
        P.w1ninv = h_mulmod(w1, P.ninv, q);
        P.w1ninv_s = h_shoup(P.w1ninv, q);
        P.ninv_d = (double)P.ninv;     P.ninv_q = (double)((long double)P.ninv / (long double)q);
        P.w1ninv_d = (double)P.w1ninv; P.w1ninv_q = (double)((long double)P.w1ninv / (long double)q);
    }
}

// Sequential powers psi^0..psi^(N-1), scattered to bit-reversed positions.
static inline u64 d_bits(double d) { u64 b; std::memcpy(&b, &d, 8); return b; }
// FP64 policy: {w, fl(w/q)} as bit patterns (x87 long double carries 64 bits, so the quotient is rounded once more to
// double: relative error <= 2^-53 (1 + 2^-11), inside fp_mulmod's budget)
static inline ulonglong2 fp_entry(u64 w, u64 q) { return make_ulonglong2(d_bits((double)w), d_bits((double)((long double)w / (long double)q))); }
static void fill_tables(ulonglong2* fwd, ulonglong2* inv, u64 q, u32 logN, u64 psi, bool f64) {
    const u64 N = 1ull << logN;
    const u64 psi_inv = h_powmod(psi, 2 * N - 1, q);
    u64 pw = 1, pwi = 1;
    for (u64 e = 0; e < N; ++e) {
        const u32 k = h_brv((u32)e, logN);
        fwd[k] = f64 ? fp_entry(pw, q) : make_ulonglong2(pw, h_shoup(pw, q));
        inv[k] = f64 ? fp_entry(pwi, q) : make_ulonglong2(pwi, h_shoup(pwi, q));
        pw = h_mulmod(pw, psi, q);
        pwi = h_mulmod(pwi, psi_inv, q);
    }
}

// t0[t / 32][j][t % 32] = std[(N >> (b+1)) + t * 2^(3-b) + hi],  j = (2^(3-b) - 1) + hi,  b = 0..3  (ntt_core.cuh, TwT0)
static void fill_t0(ulonglong2* t0, const ulonglong2* std_tab, u32 logN) {
    const u64 N = 1ull << logN, T = N >> 4;
    for (u32 b = 0; b < 4; ++b) {
        const u32 cnt = 1u << (3 - b);
        for (u32 hi = 0; hi < cnt; ++hi) {
            const u32 j = (cnt - 1) + hi;
            const ulonglong2* src = std_tab + (N >> (b + 1)) + hi;
            for (u64 t = 0; t < T; ++t) t0[(t >> 5) * kT0Block + j * 32 + (t & 31)] = src[t * cnt];
        }
    }
}

int plan_create(Plan** out, int device, u32 logN, u32 L, const u64* moduli, const u64* psi) {
    if (!out || !moduli || L == 0) return -1;
    if (logN != 0 && (logN < 4 || logN > 20 || !psi)) return -2;
    for (u32 l = 0; l < L; ++l) {
        const u64 q = moduli[l];
        if (q < 2 || q >= (1ull << 62)) return -3;
        if (logN) {
            const u64 N = 1ull << logN;
            if ((q & 1) == 0 || (q - 1) % (2 * N) != 0) return -4;           // need q = 1 mod 2N
            if (h_powmod(psi[l], N, q) != q - 1) return -5;                   // psi^N = -1
        }
    }
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess) return (int)e;
    if (device < 0 || device >= ndev) return (int)cudaErrorInvalidDevice;
    DeviceGuard guard(device);   // the caller's current device is restored on every return path
    Plan* pl = new Plan();
    pl->device = device; pl->logN = logN; pl->L = L;
    pl->moduli.assign(moduli, moduli + L);
    if (logN) pl->psi.assign(psi, psi + L);

    pl->lazy = true;
    for (u32 l = 0; l < L; ++l) { const u32 k = 64 - (u32)__builtin_clzll(moduli[l]); if (k > pl->max_bits) pl->max_bits = k; }
    for (u32 l = 0; l < L; ++l) if (moduli[l] < (1ull << 19) || moduli[l] >= (1ull << 60)) pl->lazy = false;
    pl->w32 = pl->max_bits <= 30;
    if (const char* env = std::getenv("GFHE_FORCE_HARVEY")) if (env[0] == '1') { pl->lazy = false; pl->w32 = false; }
    if (const char* env = std::getenv("GFHE_NO_W32")) if (env[0] == '1') pl->w32 = false;
    pl->f64 = !pl->w32 && pl->max_bits <= 50 && logN != 0;
    if (const char* env = std::getenv("GFHE_FORCE_HARVEY")) if (env[0] == '1') pl->f64 = false;
    if (const char* env = std::getenv("GFHE_NO_F64")) if (env[0] == '1') pl->f64 = false;
    plan_read_env(pl);
    const bool f64 = pl->f64;

    std::vector<LimbParams> hp(L);
    for (u32 l = 0; l < L; ++l) fill_limb_params(hp[l], moduli[l], logN, logN ? psi[l] : 0);
    if ((e = cudaMalloc(&pl->d_params, sizeof(LimbParams) * L)) != cudaSuccess) { plan_destroy(pl); return (int)e; }
    if ((e = cudaMemcpy(pl->d_params, hp.data(), sizeof(LimbParams) * L, cudaMemcpyHostToDevice)) != cudaSuccess) { plan_destroy(pl); return (int)e; }

    if (logN) {
        const size_t N = (size_t)1 << logN, bytes = sizeof(ulonglong2) * N * L;
        const size_t n0 = t0_entries(N), bytes0 = sizeof(ulonglong2) * n0 * L;
        ulonglong2 *hf = nullptr, *hi = nullptr, *hf0 = nullptr, *hi0 = nullptr;
        hf = (ulonglong2*)std::malloc(bytes); hi = (ulonglong2*)std::malloc(bytes);
        hf0 = (ulonglong2*)std::calloc(bytes0, 1); hi0 = (ulonglong2*)std::calloc(bytes0, 1);
        if (!hf || !hi || !hf0 || !hi0) { std::free(hf); std::free(hi); std::free(hf0); std::free(hi0); plan_destroy(pl); return -6; }
        unsigned nthr = std::thread::hardware_concurrency();
        if (nthr == 0) nthr = 1;
        if (nthr > L) nthr = L;
        if (nthr > 32) nthr = 32;
        std::vector<std::thread> pool;
        for (unsigned t = 0; t < nthr; ++t)
            pool.emplace_back([=]() {
                for (u32 l = t; l < L; l += nthr) {
                    fill_tables(hf + l * N, hi + l * N, moduli[l], logN, psi[l], f64);
                    fill_t0(hf0 + l * n0, hf + l * N, logN);
                    fill_t0(hi0 + l * n0, hi + l * N, logN);
                }
            });
        for (auto& th : pool) th.join();
        size_t up = bytes, up0 = bytes0;
        if (pl->w32) {   // 32-bit policy: compact 8-byte entries w | (w' << 32), w' = high word of the Shoup constant (ntt_core.cuh, tw_expand)
            for (ulonglong2* tab : {hf, hi, hf0, hi0}) {
                const size_t n = (tab == hf || tab == hi) ? N * L : n0 * L;
                u64* dst = reinterpret_cast<u64*>(tab);   // in place: entry i is written at byte 8i, read at byte 16i
                for (size_t i = 0; i < n; ++i) { const ulonglong2 v = tab[i]; dst[i] = (v.x & 0xffffffffull) | (v.y & 0xffffffff00000000ull); }
            }
            up = bytes / 2; up0 = bytes0 / 2;
        }
        e = cudaMalloc(&pl->d_fwd, up);
        if (e == cudaSuccess) e = cudaMalloc(&pl->d_inv, up);
        if (e == cudaSuccess) e = cudaMemcpy(pl->d_fwd, hf, up, cudaMemcpyHostToDevice);
        if (e == cudaSuccess) e = cudaMemcpy(pl->d_inv, hi, up, cudaMemcpyHostToDevice);
        if (e == cudaSuccess) e = cudaMalloc(&pl->d_fwd_t0, up0);
        if (e == cudaSuccess) e = cudaMalloc(&pl->d_inv_t0, up0);
        if (e == cudaSuccess) e = cudaMemcpy(pl->d_fwd_t0, hf0, up0, cudaMemcpyHostToDevice);
        if (e == cudaSuccess) e = cudaMemcpy(pl->d_inv_t0, hi0, up0, cudaMemcpyHostToDevice);
        std::free(hf); std::free(hi); std::free(hf0); std::free(hi0);
        if (e != cudaSuccess) { plan_destroy(pl); return (int)e; }
    }
    *out = pl;
    return 0;
}

void plan_read_env(Plan* pl) {
    if (const char* env = std::getenv("GFHE_PERSIST")) pl->persist = env[0] == '1' ? 1 : 0;
    if (const char* env = std::getenv("GFHE_PIPE")) pl->pipe = env[0] == '1' ? 1 : 0;
    if (const char* env = std::getenv("GFHE_SPLIT")) pl->split = env[0] == '1' ? 1 : 0;
}

cudaError_t pipe_sync_get(const Plan& pl, cudaStream_t st, size_t words, u32** out) {
    std::lock_guard<std::mutex> lock(pl.sync_mu);
    Plan::SyncBuf& sb = pl.sync_bufs[st];
    if (sb.words < words) {
        size_t cap = 1024;
        while (cap < words) cap *= 2;
        // the old buffer may still be read by a kernel in flight on st: cudaFree waits for the device
        if (sb.d) cudaFree(sb.d);
        sb.d = nullptr; sb.words = 0;
        cudaError_t e = cudaMalloc(&sb.d, cap * sizeof(u32));
        if (e != cudaSuccess) return e;
        if ((e = cudaMemset(sb.d, 0, cap * sizeof(u32))) != cudaSuccess) return e;
        sb.words = cap;
    }
    *out = sb.d;
    return cudaSuccess;
}

// ---- plan image: save / load of the table cache ---------------------------------------------------
namespace {
struct ImageHeader {
    char magic[8];          // "GFHEPLN" + format version
    u32 logN, L;
    u32 flags;              // bit 0 w32, bit 1 f64, bit 2 lazy: the arithmetic policy the tables are laid out for
    u32 max_bits;
    u64 table_bytes;        // bytes of d_fwd (= d_inv), all limbs
    u64 t0_bytes;           // bytes of d_fwd_t0 (= d_inv_t0)
    u64 payload_bytes;      // everything after the header
    u64 checksum;           // FNV-1a over the payload, 8 bytes at a time
};
static_assert(sizeof(ImageHeader) == 56, "plan image header layout");
const char kMagic[8] = {'G', 'F', 'H', 'E', 'P', 'L', 'N', '2'};

u64 fnv64(const unsigned char* p, size_t n) {
    u64 h = 1469598103934665603ull;
    size_t i = 0;
    for (; i + 8 <= n; i += 8) { u64 w; std::memcpy(&w, p + i, 8); h = (h ^ w) * 1099511628211ull; }
    for (; i < n; ++i) h = (h ^ p[i]) * 1099511628211ull;
    return h;
}
size_t table_bytes_of(const Plan& pl) { return pl.logN ? (sizeof(ulonglong2) << pl.logN) * pl.L / (pl.w32 ? 2 : 1) : 0; }
size_t t0_bytes_of(const Plan& pl) { return pl.logN ? sizeof(ulonglong2) * t0_entries((size_t)1 << pl.logN) * pl.L / (pl.w32 ? 2 : 1) : 0; }
}  // namespace

size_t plan_image_bytes(const Plan& pl) {
    return sizeof(ImageHeader) + sizeof(u64) * pl.L * 2 + sizeof(LimbParams) * pl.L + 2 * table_bytes_of(pl) + 2 * t0_bytes_of(pl);
}

int plan_export(const Plan& pl, void* buf, size_t bytes) {
    if (!buf) return -1;
    const size_t need = plan_image_bytes(pl);
    if (bytes < need) return -8;
    DeviceGuard guard(pl.device);
    ImageHeader h{};
    std::memcpy(h.magic, kMagic, 8);
    h.logN = pl.logN; h.L = pl.L; h.max_bits = pl.max_bits;
    h.flags = (pl.w32 ? 1u : 0u) | (pl.f64 ? 2u : 0u) | (pl.lazy ? 4u : 0u);
    h.table_bytes = table_bytes_of(pl); h.t0_bytes = t0_bytes_of(pl);
    h.payload_bytes = need - sizeof(ImageHeader);
    unsigned char* p = static_cast<unsigned char*>(buf) + sizeof(ImageHeader);
    std::memcpy(p, pl.moduli.data(), sizeof(u64) * pl.L); p += sizeof(u64) * pl.L;
    if (pl.logN) std::memcpy(p, pl.psi.data(), sizeof(u64) * pl.L); else std::memset(p, 0, sizeof(u64) * pl.L);
    p += sizeof(u64) * pl.L;
    cudaError_t e = cudaMemcpy(p, pl.d_params, sizeof(LimbParams) * pl.L, cudaMemcpyDeviceToHost); p += sizeof(LimbParams) * pl.L;
    for (const ulonglong2* t : {pl.d_fwd, pl.d_inv}) { if (e == cudaSuccess && h.table_bytes) e = cudaMemcpy(p, t, h.table_bytes, cudaMemcpyDeviceToHost); p += h.table_bytes; }
    for (const ulonglong2* t : {pl.d_fwd_t0, pl.d_inv_t0}) { if (e == cudaSuccess && h.t0_bytes) e = cudaMemcpy(p, t, h.t0_bytes, cudaMemcpyDeviceToHost); p += h.t0_bytes; }
    if (e != cudaSuccess) return (int)e;
    h.checksum = fnv64(static_cast<unsigned char*>(buf) + sizeof(ImageHeader), h.payload_bytes);
    std::memcpy(buf, &h, sizeof h);
    return 0;
}

int plan_import(Plan** out, int device, const void* buf, size_t bytes) {
    if (!out || !buf) return -1;
    if (bytes < sizeof(ImageHeader)) return -14;
    ImageHeader h;
    std::memcpy(&h, buf, sizeof h);
    if (std::memcmp(h.magic, kMagic, 8) != 0 || h.L == 0 || (h.logN != 0 && (h.logN < 4 || h.logN > 20))) return -14;
    Plan probe; probe.logN = h.logN; probe.L = h.L; probe.w32 = (h.flags & 1u) != 0;
    if (h.table_bytes != table_bytes_of(probe) || h.t0_bytes != t0_bytes_of(probe) || sizeof(ImageHeader) + h.payload_bytes != plan_image_bytes(probe) ||
        bytes < plan_image_bytes(probe))
        return -14;
    const unsigned char* p = static_cast<const unsigned char*>(buf) + sizeof(ImageHeader);
    if (fnv64(p, h.payload_bytes) != h.checksum) return -14;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess) return (int)e;
    if (device < 0 || device >= ndev) return (int)cudaErrorInvalidDevice;
    DeviceGuard guard(device);
    Plan* pl = new Plan();
    pl->device = device; pl->logN = h.logN; pl->L = h.L; pl->max_bits = h.max_bits;
    pl->w32 = (h.flags & 1u) != 0; pl->f64 = (h.flags & 2u) != 0; pl->lazy = (h.flags & 4u) != 0;   // the policy the tables were built for
    plan_read_env(pl);
    pl->moduli.resize(h.L); std::memcpy(pl->moduli.data(), p, sizeof(u64) * h.L); p += sizeof(u64) * h.L;
    if (h.logN) { pl->psi.resize(h.L); std::memcpy(pl->psi.data(), p, sizeof(u64) * h.L); }
    p += sizeof(u64) * h.L;
    for (u32 l = 0; l < h.L; ++l) {   // the constants must belong to the moduli the header names
        LimbParams lp; std::memcpy(&lp, p + sizeof(LimbParams) * l, sizeof lp);
        if (lp.q != pl->moduli[l] || lp.q < 2 || lp.q >= (1ull << 62)) { delete pl; return -14; }
    }
    if ((e = cudaMalloc(&pl->d_params, sizeof(LimbParams) * h.L)) == cudaSuccess) e = cudaMemcpy(pl->d_params, p, sizeof(LimbParams) * h.L, cudaMemcpyHostToDevice);
    p += sizeof(LimbParams) * h.L;
    for (ulonglong2** t : {&pl->d_fwd, &pl->d_inv}) {
        if (e == cudaSuccess && h.table_bytes) { if ((e = cudaMalloc(t, h.table_bytes)) == cudaSuccess) e = cudaMemcpy(*t, p, h.table_bytes, cudaMemcpyHostToDevice); }
        p += h.table_bytes;
    }
    for (ulonglong2** t : {&pl->d_fwd_t0, &pl->d_inv_t0}) {
        if (e == cudaSuccess && h.t0_bytes) { if ((e = cudaMalloc(t, h.t0_bytes)) == cudaSuccess) e = cudaMemcpy(*t, p, h.t0_bytes, cudaMemcpyHostToDevice); }
        p += h.t0_bytes;
    }
    if (e != cudaSuccess) { plan_destroy(pl); return (int)e; }
    *out = pl;
    return 0;
}

// ---- policy dispatch ---------------------------------------------------------------------------
extern template struct Launcher<Harvey>;
extern template struct Launcher<Lazy60>;
extern template struct Launcher<Shoup32>;
extern template struct Launcher<Fp50>;
extern template struct Launcher<Shoup32P>;

// one policy's launcher for a limb range, single stream
static cudaError_t do_ntt(const Plan& pl, bool inv, u32 limb0, u32 nl, const u64* in, u64* out, u32 B, cudaStream_t st) {
    if (inv) {
        if (pl.w32) return Launcher<Shoup32>::ntt_inv(pl, limb0, nl, in, out, B, st);
        if (pl.f64) return Launcher<Fp50>::ntt_inv(pl, limb0, nl, in, out, B, st);
        return pl.lazy ? Launcher<Lazy60>::ntt_inv(pl, limb0, nl, in, out, B, st) : Launcher<Harvey>::ntt_inv(pl, limb0, nl, in, out, B, st);
    }
    if (pl.w32) return Launcher<Shoup32>::ntt_fwd(pl, limb0, nl, in, out, B, st);
    if (pl.f64) return Launcher<Fp50>::ntt_fwd(pl, limb0, nl, in, out, B, st);
    return pl.lazy ? Launcher<Lazy60>::ntt_fwd(pl, limb0, nl, in, out, B, st) : Launcher<Harvey>::ntt_fwd(pl, limb0, nl, in, out, B, st);
}
static cudaError_t do_poly_mul(const Plan& pl, u32 limb0, u32 nl, const u64* a, const u64* b, u64* c, u64* ws, u32 B, cudaStream_t st) {
    if (pl.w32) return Launcher<Shoup32>::poly_mul(pl, limb0, nl, a, b, c, ws, B, st);
    if (pl.f64) return Launcher<Fp50>::poly_mul(pl, limb0, nl, a, b, c, ws, B, st);
    return pl.lazy ? Launcher<Lazy60>::poly_mul(pl, limb0, nl, a, b, c, ws, B, st)
                   : Launcher<Harvey>::poly_mul(pl, limb0, nl, a, b, c, ws, B, st);
}
static cudaError_t do_poly_mul_pre(const Plan& pl, u32 limb0, u32 nl, const u64* a, const u64* b_ntt, u64* c, u32 B, cudaStream_t st) {
    if (pl.w32) return Launcher<Shoup32>::poly_mul_pre(pl, limb0, nl, a, b_ntt, c, B, st);
    if (pl.f64) return Launcher<Fp50>::poly_mul_pre(pl, limb0, nl, a, b_ntt, c, B, st);
    return pl.lazy ? Launcher<Lazy60>::poly_mul_pre(pl, limb0, nl, a, b_ntt, c, B, st)
                   : Launcher<Harvey>::poly_mul_pre(pl, limb0, nl, a, b_ntt, c, B, st);
}

// ---- two-stream split of small jobs ---------------------------------------------------------------
// A two-pass transform is one grid per pass and the product three; with a few hundred to a few thousand tiles (a limb shard
// of config 4 on 8 GPUs is 1024 tiles per grid on 296-592 resident CTAs) the last, partly filled wave of every grid leaves
// SMs idle.  Such a job is cut into two limb halves that run on the caller's stream and on a stream of the plan's own,
// forked and joined with events, so one half's tails are filled by the other half's kernels.  Measured on the 8-GPU shard
// [4,16,65536] (profiles/r02_run7_chunk_probe_shard.txt): poly-mul 0.1402 -> 0.1317 ms, NTT 0.0515 -> 0.0472 ms; at the
// full [32,16,65536] the grids are 14-28 waves deep and the split gains nothing, so it is applied to small jobs only.
// GFHE_SPLIT=0 / 1 at plan creation: never / whenever there are two limbs.
static bool want_split(const Plan& pl, u32 nl, u32 B) {
    if (pl.split == 0 || pl.logN <= 12 || (nl < 2 && B < 2) || pl.pipe == 1) return false;
    if (pl.split == 1) return true;
    const u64 tiles = ((u64)nl * B) << (pl.logN - 12);
    if (tiles > 4096) return pl.w32 && nl >= 4;     // large jobs: the L2-sized chunk schedule, 32-bit words only (split_run)
    return tiles >= 512;
}
// run(first limb, limbs, batch of this half, offset of the half in polynomials, stream).  Two or more limbs: cut between limbs;
// a single limb: cut its batch.
template <class F>
static cudaError_t split_run(const Plan& pl, u32 limb0, u32 nl, u32 B, cudaStream_t st, F run) {
    if (!want_split(pl, nl, B)) return run(limb0, nl, B, (size_t)0, st);
    std::lock_guard<std::mutex> lock(pl.sync_mu);   // the fork / join events are the plan's: one enqueue sequence at a time
    cudaError_t e;
    if (!pl.aux_stream) {
        if ((e = cudaStreamCreateWithFlags(&pl.aux_stream, cudaStreamNonBlocking))) return e;
        if ((e = cudaEventCreateWithFlags(&pl.ev_fork, cudaEventDisableTiming))) return e;
        if ((e = cudaEventCreateWithFlags(&pl.ev_join, cudaEventDisableTiming))) return e;
    }
    if ((e = cudaEventRecord(pl.ev_fork, st))) return e;
    if ((e = cudaStreamWaitEvent(pl.aux_stream, pl.ev_fork, 0))) return e;
    const u64 limb_tiles = (u64)B << (pl.logN - 12);
    if (pl.split < 0 && limb_tiles * nl > 4096) {
        // large job on 32-bit words (want_split): chunks of about 2048 tiles alternate between the two streams, so a chunk's
        // inter-pass words are re-read while they are still in L2 and each stream's launch tails are filled by the other's
        // kernels.  Measured at [32,16,65536] (profiles/r02_run3_chunk_probe.txt, 8-limb chunks): 30-bit product 0.463 ->
        // 0.442 ms, packed 0.359 -> 0.339 ms, NTT -3 %; nothing for 60-bit limbs (issue-bound), which is why they are excluded.
        u32 cl = (u32)(2048 / (limb_tiles ? limb_tiles : 1));
        if (cl < 1) cl = 1;
        u32 k = 0;
        for (u32 l = 0; l < nl && e == cudaSuccess; l += cl, ++k) {
            const u32 n = nl - l < cl ? nl - l : cl;
            e = run(limb0 + l, n, B, (size_t)l * B, (k & 1) ? pl.aux_stream : st);
        }
    } else if (nl >= 2) {
        const u32 h = nl / 2;
        if ((e = run(limb0, h, B, (size_t)0, st))) return e;
        e = run(limb0 + h, nl - h, B, (size_t)h * B, pl.aux_stream);
    } else {
        const u32 h = B / 2;
        if ((e = run(limb0, 1u, h, (size_t)0, st))) return e;
        e = run(limb0, 1u, B - h, (size_t)h, pl.aux_stream);
    }
    // join even after a failed launch: the caller's stream must not run ahead of what was enqueued on the other one
    cudaError_t e2 = cudaEventRecord(pl.ev_join, pl.aux_stream);
    if (e2 == cudaSuccess) e2 = cudaStreamWaitEvent(st, pl.ev_join, 0);
    return e ? e : e2;
}

cudaError_t launch_ntt_fwd(const Plan& pl, u32 limb0, u32 nl, const u64* in, u64* out, u32 B, cudaStream_t st) {
    const u32 lg = pl.logN;
    return split_run(pl, limb0, nl, B, st, [&](u32 l0, u32 n, u32 b, size_t off, cudaStream_t s) { return do_ntt(pl, false, l0, n, in + (off << lg), out + (off << lg), b, s); });
}
cudaError_t launch_ntt_inv(const Plan& pl, u32 limb0, u32 nl, const u64* in, u64* out, u32 B, cudaStream_t st) {
    const u32 lg = pl.logN;
    return split_run(pl, limb0, nl, B, st, [&](u32 l0, u32 n, u32 b, size_t off, cudaStream_t s) { return do_ntt(pl, true, l0, n, in + (off << lg), out + (off << lg), b, s); });
}
cudaError_t launch_poly_mul(const Plan& pl, u32 limb0, u32 nl, const u64* a, const u64* b, u64* c, u64* ws, u32 B, cudaStream_t st) {
    const u32 lg = pl.logN;
    return split_run(pl, limb0, nl, B, st, [&](u32 l0, u32 n, u32 bb, size_t off, cudaStream_t s) {
        return do_poly_mul(pl, l0, n, a + (off << lg), b + (off << lg), c + (off << lg), ws ? ws + (off << lg) : nullptr, bb, s); });
}
cudaError_t launch_poly_mul_pre(const Plan& pl, u32 limb0, u32 nl, const u64* a, const u64* b_ntt, u64* c, u32 B, cudaStream_t st) {
    const u32 lg = pl.logN;
    return split_run(pl, limb0, nl, B, st, [&](u32 l0, u32 n, u32 bb, size_t off, cudaStream_t s) { return do_poly_mul_pre(pl, l0, n, a + (off << lg), b_ntt + (off << lg), c + (off << lg), bb, s); });
}
// packed layout: only plans on the 32-bit policy have it (every modulus below 2^30)
cudaError_t launch_ntt_fwd_u32(const Plan& pl, u32 limb0, u32 nl, const u32* in, u32* out, u32 B, cudaStream_t st) {
    if (!pl.w32) return cudaErrorNotSupported;
    const u32 lg = pl.logN;
    return split_run(pl, limb0, nl, B, st, [&](u32 l0, u32 n, u32 b, size_t off, cudaStream_t s) { return Launcher<Shoup32P>::ntt_fwd(pl, l0, n, in + (off << lg), out + (off << lg), b, s); });
}
cudaError_t launch_ntt_inv_u32(const Plan& pl, u32 limb0, u32 nl, const u32* in, u32* out, u32 B, cudaStream_t st) {
    if (!pl.w32) return cudaErrorNotSupported;
    const u32 lg = pl.logN;
    return split_run(pl, limb0, nl, B, st, [&](u32 l0, u32 n, u32 b, size_t off, cudaStream_t s) { return Launcher<Shoup32P>::ntt_inv(pl, l0, n, in + (off << lg), out + (off << lg), b, s); });
}
cudaError_t launch_poly_mul_u32(const Plan& pl, u32 limb0, u32 nl, const u32* a, const u32* b, u32* c, u32* ws, u32 B, cudaStream_t st) {
    if (!pl.w32) return cudaErrorNotSupported;
    const u32 lg = pl.logN;
    return split_run(pl, limb0, nl, B, st, [&](u32 l0, u32 n, u32 bb, size_t off, cudaStream_t s) {
        return Launcher<Shoup32P>::poly_mul(pl, l0, n, a + (off << lg), b + (off << lg), c + (off << lg), ws ? ws + (off << lg) : nullptr, bb, s); });
}
cudaError_t launch_poly_mul_pre_u32(const Plan& pl, u32 limb0, u32 nl, const u32* a, const u32* b_ntt, u32* c, u32 B, cudaStream_t st) {
    if (!pl.w32) return cudaErrorNotSupported;
    const u32 lg = pl.logN;
    return split_run(pl, limb0, nl, B, st, [&](u32 l0, u32 n, u32 bb, size_t off, cudaStream_t s) { return Launcher<Shoup32P>::poly_mul_pre(pl, l0, n, a + (off << lg), b_ntt + (off << lg), c + (off << lg), bb, s); });
}
size_t poly_mul_workspace_bytes(const Plan& pl, u32 nl, u32 B) {
    if (pl.logN <= 12) return 0;
    return ((size_t)nl * B * sizeof(u64)) << pl.logN;   // one operand-sized scratch
}

// ---- product on the library's own scratch: an entry point that allocates nothing -----------------------------------
// For callers that cannot call cudaMalloc (stream capture, serving loops, harnesses that forbid it): gfhe_poly_mul's
// operand-sized workspace lives in a file-scope device array instead of a caller buffer.  Every GPU whose context loads the
// library holds kStaticWsBytes of it until the process exits.  In host code `g_static_ws` is only the host-side stand-in of
// the symbol: the device address is asked of the runtime once per device, with that device current, and that pointer is
// what the kernels get.  The product writes every scratch word it later reads (the column pass of b), so nothing has to
// start at zero; all calls on one GPU share the array, so two calls that overlap on the GPU race.
static __device__ __align__(256) unsigned char g_static_ws[kStaticWsBytes];
static std::mutex g_static_mu;
static void* g_static_addr[64];          // per device ordinal, resolved on first use

cudaError_t launch_poly_mul_static(const Plan& pl, u32 limb0, u32 nl, const u64* a, const u64* b, u64* c, u32 B, cudaStream_t st) {
    if (pl.device < 0 || pl.device >= 64) return cudaErrorInvalidDevice;
    void* ws = nullptr;
    {
        std::lock_guard<std::mutex> lock(g_static_mu);
        if (!g_static_addr[pl.device]) {    // the caller holds a DeviceGuard on pl.device
            void* p = nullptr;
            if (cudaError_t e = cudaGetSymbolAddress(&p, g_static_ws)) return e;
            g_static_addr[pl.device] = p;
        }
        ws = g_static_addr[pl.device];
    }
    // one stream, no two-stream split (its stream and events are created on first use) and no pipelined kernel (its
    // synchronisation words are allocated per stream): straight to the policy's three launches
    return do_poly_mul(pl, limb0, nl, a, b, c, (u64*)ws, B, st);
}

void host_pipe_destroy(Plan* p);   // capi.cu

int plan_destroy(Plan* p) {
    if (!p) return 0;
    DeviceGuard guard(p->device);
    host_pipe_destroy(p);
    if (p->prof) {
        for (auto& r : p->prof->recs) { cudaEventDestroy(r.a); cudaEventDestroy(r.b); }
        for (auto e : p->prof->pool) cudaEventDestroy(e);
        delete p->prof;
    }
    for (auto& kv : p->sync_bufs) if (kv.second.d) cudaFree(kv.second.d);
    if (p->ev_fork) cudaEventDestroy(p->ev_fork);
    if (p->ev_join) cudaEventDestroy(p->ev_join);
    if (p->aux_stream) cudaStreamDestroy(p->aux_stream);
    if (p->d_params) cudaFree(p->d_params);
    if (p->d_fwd) cudaFree(p->d_fwd);
    if (p->d_inv) cudaFree(p->d_inv);
    if (p->d_fwd_t0) cudaFree(p->d_fwd_t0);
    if (p->d_inv_t0) cudaFree(p->d_inv_t0);
    delete p;
    return 0;
}

}  // namespace gfhe
