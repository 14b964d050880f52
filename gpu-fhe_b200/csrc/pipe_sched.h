// Ticket schedule of the pipelined kernels (ntt_pipe.inl): plain host/device code, no kernels, so the C ABI's schedule
// check (gfhe_debug_pipe_schedule, tests/test_pipe_schedule.py) exercises exactly what the kernels decode.
#pragma once
#include "modarith.cuh"

namespace gfhe {

struct PipeSched {
    u32 nph;          // passes: 2 (transform) or 3 (product)
    u32 npoly, lag, total;
    u32 n[3];         // tiles per polynomial in each pass
    u32 nr;           // ranges of steps with a constant set of active passes
    u32 tb[7];        // first ticket of range r; tb[nr] = total
    u32 s0[6];        // first step of range r
    u32 cnt[6];       // tickets per step in range r
    u32 mask[6];      // active passes of range r (bit k = pass k, working on polynomial step - k*lag)
};

// ticket -> (pass, polynomial, tile of the pass); shared by the kernels and the host-side schedule check
__host__ __device__ inline void pipe_decode(const PipeSched& sc, u32 t, u32* pass_out, u32* poly_out, u32* tile_out) {
    u32 r = 0;
    while (r + 1 < sc.nr && t >= sc.tb[r + 1]) ++r;
    const u32 d = t - sc.tb[r];
    const u32 step = sc.s0[r] + d / sc.cnt[r];
    u32 w = d % sc.cnt[r];
    u32 pass = 0;
    for (; pass < sc.nph; ++pass) {
        if ((sc.mask[r] >> pass) & 1u) {
            if (w < sc.n[pass]) break;
            w -= sc.n[pass];
        }
    }
    *pass_out = pass; *poly_out = step - pass * sc.lag; *tile_out = w;
}

// Ticket schedule: step s works on polynomial s - k*lag in pass k; ranges of steps with the same set of live passes.
inline PipeSched make_pipe_sched(u32 nph, const u32* n, u32 npoly, u32 resident) {
    PipeSched sc{};
    sc.nph = nph; sc.npoly = npoly;
    u32 z = 0;
    for (u32 k = 0; k < nph; ++k) { sc.n[k] = n[k]; z += n[k]; }
    // a pass's tiles should have left the SMs when their consumers are drawn: lag * (tickets per step) >= resident CTAs + a margin
    u32 lag = (resident + resident / 4 + z - 1) / z;
    if (lag < 1) lag = 1;
    if (lag > npoly) lag = npoly;
    sc.lag = lag;
    u32 cuts[8], nc = 0;
    for (u32 k = 0; k < nph; ++k) { cuts[nc++] = k * lag; cuts[nc++] = npoly + k * lag; }
    for (u32 i = 0; i < nc; ++i) for (u32 j = i + 1; j < nc; ++j) if (cuts[j] < cuts[i]) { const u32 t = cuts[i]; cuts[i] = cuts[j]; cuts[j] = t; }
    u32 tb = 0;
    for (u32 i = 0; i + 1 < nc; ++i) {
        const u32 lo = cuts[i], hi = cuts[i + 1];
        if (hi == lo) continue;
        u32 mask = 0, cnt = 0;
        for (u32 k = 0; k < nph; ++k) if (lo >= k * lag && lo < npoly + k * lag) { mask |= 1u << k; cnt += n[k]; }
        if (!cnt) continue;
        sc.tb[sc.nr] = tb; sc.s0[sc.nr] = lo; sc.cnt[sc.nr] = cnt; sc.mask[sc.nr] = mask; ++sc.nr;
        tb += cnt * (hi - lo);
    }
    sc.tb[sc.nr] = tb;
    sc.total = tb;
    return sc;
}

}  // namespace gfhe
