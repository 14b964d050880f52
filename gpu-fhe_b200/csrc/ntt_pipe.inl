// Pipelined two-pass transforms: ONE launch, intermediate kept in L2 (included by ntt_kernels.inl).
//
// A transform of N > 4096 points is a column pass and a row pass over 4096-residue tiles; the fused product is three
// passes.  Launched as separate kernels, every pass streams its whole operand set through HBM (2-3x the algorithmic
// bytes: profiles/ncu_traffic.json, 2342 MB per config-4 step against 805 MB).  Here one persistent kernel runs all the
// passes: CTAs draw tickets from a global counter, and the ticket order interleaves the passes with a lag of a few
// polynomials --
//
//     ... col(p + 2*lag) | row(p + lag) | col_inv(p) | col(p + 2*lag + 1) | row(p + lag + 1) | col_inv(p + 1) ...
//
// -- so what a pass writes is consumed a few hundred tickets later, out of B200's 126 MB L2, and HBM sees the operands
// once in and the result once out.  A consumer tile waits (one thread, acquire loads) until every producer tile of its
// polynomial has been counted as done.  Tickets are handed out in order and a ticket's producers all have smaller
// numbers, so every tile anybody waits for is already running on a resident CTA (or finished): the wait cannot deadlock
// whatever the grid size, and dynamic tickets also remove the wave quantisation of 3 separate grids.
//
// Memory ordering: producer threads store, __threadfence(), barrier, then thread 0 bumps the polynomial's counter; the
// consumer's thread 0 spins on ld.acquire.gpu, the CTA barrier hands the ordering to the other threads, and they read the
// producer's words through L2 only (ld.global.cg / TMA) -- this SM's L1 may hold what those addresses held before.
//
// The synchronisation words live in a per-(plan, stream) buffer that the last CTA of every launch leaves zeroed.

// sync[0] = ticket counter, sync[1] = CTAs that have left, sync[2 + k*npoly + p] = finished tiles of pass k, polynomial p
__device__ __forceinline__ u32 ld_acquire_gpu(const u32* p) {
    u32 v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// thread 0: decode ticket t and wait for the tiles it consumes; then draw the NEXT ticket at once -- its L2 round trip runs
// under this tile's work instead of between two tiles (the value is first looked at on the next trip)
// PREFETCH = false (the product kernel, whose fused row pass has no register to spare for a value that lives across the
// tile): the ticket is drawn here instead, `t` is ignored.
template <bool PREFETCH>
__device__ __forceinline__ u32 pipe_next(const PipeSched& sc, u32* sync, volatile u32* job, u32 t) {
    if constexpr (!PREFETCH) t = atomicAdd(sync, 1u);
    if (t >= sc.total) { job[0] = sc.nph; return t; }
    u32 pass, poly, w;
    pipe_decode(sc, t, &pass, &poly, &w);
    job[0] = pass; job[1] = poly; job[2] = w;
    if (pass > 0) {
        const u32* flag = sync + 2 + (size_t)(pass - 1) * sc.npoly + poly;
        const u32 need = sc.n[pass - 1];
        // bounded: producers are running CTAs, so the wait is microseconds; a schedule bug must fail the launch, not hang the GPU
        u32 polls = 0;
        while (ld_acquire_gpu(flag) < need) {
            __nanosleep(64);
            if (++polls > (1u << 23)) asm volatile("trap;");
        }
        asm volatile("fence.proxy.async;" ::: "memory");   // the producers used generic stores; this thread may read them by TMA next
    }
    if constexpr (PREFETCH) return atomicAdd(sync, 1u);
    else return 0u;
}

// after a tile: the barrier orders every thread's global stores before thread 0's fence (cumulativity -- the pattern of a
// cooperative-groups grid barrier), and thread 0 alone releases them and counts the tile
// (pass and polynomial are read back from the job words rather than kept in registers across the tile body)
__device__ __forceinline__ void pipe_done(const PipeSched& sc, u32* sync, const volatile u32* job) {
    __syncthreads();
    if (threadIdx.x == 0) {
        const u32 pass = job[0], poly = job[1];
        if (pass + 1 < sc.nph) {
            __threadfence();
            atomicAdd(sync + 2 + (size_t)pass * sc.npoly + poly, 1u);
        }
    }
}

// the last CTA to leave zeroes the words for the next launch on this stream
__device__ __forceinline__ void pipe_exit(const PipeSched& sc, u32* sync, volatile u32* job) {
    __syncthreads();   // every thread has read the job that ended its loop
    if (threadIdx.x == 0) {
        __threadfence();
        job[0] = atomicAdd(sync + 1, 1u);
    }
    __syncthreads();
    if (job[0] == gridDim.x - 1) {
        const u32 words = 2 + (sc.nph - 1) * sc.npoly;
        for (u32 i = threadIdx.x; i < words; i += blockDim.x) sync[i] = 0;
    }
}

template <class A> struct PipeOcc {   // resident CTAs per SM the transform pipelines are compiled for (the row pass sets the register budget)
    static constexpr int kNtt = RowOcc<A, 12>::kB;
};

// forward: column pass (in -> out), row pass (out -> out)
template <class A, int LOGC, int S>
__global__ void __launch_bounds__(256, PipeOcc<A>::kNtt)
k_pipe_ntt_fwd(const typename A::store* in, typename A::store* out, const Tables tabs, u32 B, u32 limb0, const PipeSched sc, u32* sync) {
    typedef typename A::word W;
    constexpr u32 LOGN = S + 12 - LOGC, LT = LOGN - 12;
    __shared__ __align__(16) W sm[padded_words(12)];
    __shared__ typename A::tw_t tws[256];
    __shared__ u32 job[3];
    u32 ticket = threadIdx.x == 0 ? atomicAdd(sync, 1u) : 0u;
    for (;;) {
        if (threadIdx.x == 0) ticket = pipe_next<true>(sc, sync, job, ticket);
        __syncthreads();
        const u32 pass = job[0], flat = (job[1] << LT) + job[2];
        if (pass >= 2) break;
        if (pass == 0) col_fwd_body<A, LOGC, S>(in, out, tabs, B, limb0, flat, sm, tws);
        else row_fwd_body<A, S, 12, true>(out, out, tabs, LOGN, B, limb0, flat, sm);
        pipe_done(sc, sync, job);
    }
    pipe_exit(sc, sync, job);
}

// inverse: row pass (in -> out), column pass (out -> out)
template <class A, int LOGC, int S>
__global__ void __launch_bounds__(256, PipeOcc<A>::kNtt)
k_pipe_ntt_inv(const typename A::store* in, typename A::store* out, const Tables tabs, u32 B, u32 limb0, const PipeSched sc, u32* sync) {
    typedef typename A::word W;
    constexpr u32 LOGN = S + 12 - LOGC, LT = LOGN - 12;
    __shared__ __align__(16) W sm[padded_words(12)];
    __shared__ typename A::tw_t tws[256];
    __shared__ u32 job[3];
    u32 ticket = threadIdx.x == 0 ? atomicAdd(sync, 1u) : 0u;
    for (;;) {
        if (threadIdx.x == 0) ticket = pipe_next<true>(sc, sync, job, ticket);
        __syncthreads();
        const u32 pass = job[0], flat = (job[1] << LT) + job[2];
        if (pass >= 2) break;
        if (pass == 0) row_inv_body<A, S, 12, false>(in, out, tabs, LOGN, B, limb0, flat, sm);
        else col_inv_body<A, LOGC, S, true>(out, out, tabs, B, limb0, flat, sm, tws);
        pipe_done(sc, sync, job);
    }
    pipe_exit(sc, sync, job);
}

// product: column pass of a (-> c) and b (-> ws), fused row pass (c, ws -> c), inverse column pass (c -> c).
// BNTT: b is a cached NTT(b): no column pass for it, the row pass reads it as it is.
template <class A, int LOGC, int S, bool BNTT>
__global__ void __launch_bounds__(256, MulSmem<A, 12, false>::kMinB)
k_pipe_polymul(const typename A::store* a, const typename A::store* b, typename A::store* c, typename A::store* ws, const Tables tf, const Tables ti,
               u32 B, u32 limb0, const PipeSched sc, u32* sync) {
    typedef typename A::word W;
    typedef typename A::tw_t E;
    using Lay = MulSmem<A, 12, false>;
    constexpr u32 LOGN = S + 12 - LOGC, LT = LOGN - 12, TP = 1u << LT;
    extern __shared__ __align__(16) unsigned char smem[];
    W* sm0 = reinterpret_cast<W*>(smem);                               // exchange buffer of every pass
    E* tws = reinterpret_cast<E*>(smem + Lay::kBytes);                  // staged twiddles of the column passes
    u64* mbar = reinterpret_cast<u64*>(smem + Lay::kOffBar);
    __shared__ u32 job[4];   // job[3]: phase of the TMA barrier = fused row tiles this CTA has run, mod 2 (kept out of the registers)
    if (threadIdx.x == 0) {
        mbar_init(mbar);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        job[3] = 0;
    }
#pragma unroll 1
    for (;;) {
        if (threadIdx.x == 0) pipe_next<false>(sc, sync, job, 0u);
        __syncthreads();
        const u32 pass = job[0], poly = job[1], w = job[2];
        if (pass >= 3) break;
        if (pass == 0) {
            if (BNTT || w < TP) col_fwd_body<A, LOGC, S>(a, c, tf, B, limb0, (poly << LT) + w, sm0, tws);
            else col_fwd_body<A, LOGC, S>(b, ws, tf, B, limb0, (poly << LT) + (w - TP), sm0, tws);
        } else if (pass == 1) {
            row_polymul_body<A, S, 12, false, BNTT, true>(c, BNTT ? b : ws, c, tf, ti, LOGN, B, limb0, (poly << LT) + w, smem, job + 3);
            __syncthreads();
            if (threadIdx.x == 0) job[3] ^= 1u;
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // b's landing buffer was read by generic loads; a TMA copy writes it next
        } else {
            col_inv_body<A, LOGC, S, true>(c, c, ti, B, limb0, (poly << LT) + w, sm0, tws);
        }
        pipe_done(sc, sync, job);
    }
    pipe_exit(sc, sync, job);
}

// ---- host side ---------------------------------------------------------------------------------------------------------
// resident CTAs of a pipelined kernel on the plan's device (cached per instantiation and device)
template <class K>
static cudaError_t pipe_resident(K kernel, size_t smem, const Plan& pl, std::atomic<int>* slots, int* out) {
    const int dev = pl.device & 63;
    int n = slots[dev].load(std::memory_order_acquire);
    if (n == 0) {
        cudaError_t e;
        if (smem > 48 * 1024 && (e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)) != cudaSuccess) return e;
        int sms = 148, per_sm = 1;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, pl.device);
        if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, 256, smem)) != cudaSuccess) return e;
        n = sms * (per_sm > 0 ? per_sm : 1);
        slots[dev].store(n, std::memory_order_release);
    }
    *out = n;
    return cudaSuccess;
}

template <class A, int LOGC, int S>
static cudaError_t launch_pipe_ntt(const Plan& pl, bool inverse, u32 limb0, const typename A::store* in, typename A::store* out, u32 B, u32 npoly, cudaStream_t st) {
    static std::atomic<int> slots_f[64], slots_i[64];
    int resident = 0;
    cudaError_t e = inverse ? pipe_resident(k_pipe_ntt_inv<A, LOGC, S>, 0, pl, slots_i, &resident) : pipe_resident(k_pipe_ntt_fwd<A, LOGC, S>, 0, pl, slots_f, &resident);
    if (e != cudaSuccess) return e;
    const u32 tp = 1u << (pl.logN - 12);
    const u32 n[3] = {tp, tp, 0};
    const PipeSched sc = make_pipe_sched(2, n, npoly, (u32)resident);
    u32* sync = nullptr;
    if ((e = pipe_sync_get(pl, st, 2 + (size_t)npoly, &sync)) != cudaSuccess) return e;
    const unsigned grid = sc.total < (u32)resident ? sc.total : (unsigned)resident;
    ProfScope ps(pl, inverse ? K_PIPE_INV : K_PIPE_FWD, st);
    if (inverse) k_pipe_ntt_inv<A, LOGC, S><<<grid, 256, 0, st>>>(in, out, inv_tables(pl), B, limb0, sc, sync);
    else k_pipe_ntt_fwd<A, LOGC, S><<<grid, 256, 0, st>>>(in, out, fwd_tables(pl), B, limb0, sc, sync);
    return cudaGetLastError();
}

template <class A, int LOGC, int S, bool BNTT>
static cudaError_t launch_pipe_polymul(const Plan& pl, u32 limb0, const typename A::store* a, const typename A::store* b, typename A::store* c,
                                       typename A::store* ws, u32 B, u32 npoly, cudaStream_t st) {
    static std::atomic<int> slots[64];
    const size_t smem = MulSmem<A, 12, false>::kBytes + 256 * sizeof(typename A::tw_t);
    int resident = 0;
    cudaError_t e = pipe_resident(k_pipe_polymul<A, LOGC, S, BNTT>, smem, pl, slots, &resident);
    if (e != cudaSuccess) return e;
    const u32 tp = 1u << (pl.logN - 12);
    const u32 n[3] = {BNTT ? tp : 2 * tp, tp, tp};
    const PipeSched sc = make_pipe_sched(3, n, npoly, (u32)resident);
    u32* sync = nullptr;
    if ((e = pipe_sync_get(pl, st, 2 + 2 * (size_t)npoly, &sync)) != cudaSuccess) return e;
    const unsigned grid = sc.total < (u32)resident ? sc.total : (unsigned)resident;
    ProfScope ps(pl, K_PIPE_MUL, st);
    k_pipe_polymul<A, LOGC, S, BNTT><<<grid, 256, smem, st>>>(a, b, c, ws, fwd_tables(pl), inv_tables(pl), B, limb0, sc, sync);
    return cudaGetLastError();
}

// dispatch on the stage split of the plan's N (same table as launch_col_fwd): LOGC = 12 - (logN - S)
#define GFHE_PIPE_SWITCH(CALL8, CALL12)                                   \
    switch ((12 - (int)(pl.logN - row_stages(pl.logN))) * 16 + (int)row_stages(pl.logN)) { \
        case 4 * 16 + 8: return CALL8(4);  case 5 * 16 + 8: return CALL8(5);  case 6 * 16 + 8: return CALL8(6);  case 7 * 16 + 8: return CALL8(7);  \
        case 4 * 16 + 12: return CALL12(4); case 5 * 16 + 12: return CALL12(5); case 6 * 16 + 12: return CALL12(6); case 7 * 16 + 12: return CALL12(7); \
        default: return cudaErrorInvalidValue;                            \
    }

template <class A>
static cudaError_t pipe_ntt(const Plan& pl, bool inverse, u32 limb0, u32 nl, const typename A::store* in, typename A::store* out, u32 B, cudaStream_t st) {
    const u32 npoly = nl * B;
#define C8(LC) launch_pipe_ntt<A, LC, 8>(pl, inverse, limb0, in, out, B, npoly, st)
#define C12(LC) launch_pipe_ntt<A, LC, 12>(pl, inverse, limb0, in, out, B, npoly, st)
    GFHE_PIPE_SWITCH(C8, C12)
#undef C8
#undef C12
}

template <class A, bool BNTT>
static cudaError_t pipe_polymul(const Plan& pl, u32 limb0, u32 nl, const typename A::store* a, const typename A::store* b, typename A::store* c,
                                typename A::store* ws, u32 B, cudaStream_t st) {
    const u32 npoly = nl * B;
#define C8(LC) launch_pipe_polymul<A, LC, 8, BNTT>(pl, limb0, a, b, c, ws, B, npoly, st)
#define C12(LC) launch_pipe_polymul<A, LC, 12, BNTT>(pl, limb0, a, b, c, ws, B, npoly, st)
    GFHE_PIPE_SWITCH(C8, C12)
#undef C8
#undef C12
}
