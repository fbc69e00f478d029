// Host-side helpers shared by the translation units that launch the marching kernels (cs_stage.cu: one stage per launch;
// cs_step_kernel.cu: one step per launch).
#pragma once

#include <cstdint>
#include <map>
#include <mutex>
#include <tuple>

#include "cs_internal.cuh"
#include "cs_cell.cuh"
#include "cs_stage_march.cuh"

namespace cs_launch {

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

template <typename R>
StageParams<R> make_params(cs_ctx* ctx, const void* const q0[3], const void* const qs[3], void* const qout[3], double a,
                           double b, double dt, int part) {
    StageParams<R> P;
    P.h0 = (const R*)q0[0], P.u0 = (const R*)q0[1], P.v0 = (const R*)q0[2];
    P.hq = (const R*)qs[0], P.uq = (const R*)qs[1], P.vq = (const R*)qs[2];
    P.ho = (R*)qout[0], P.uo = (R*)qout[1], P.vo = (R*)qout[2];
    P.topo = ctx->has_topo ? (const R*)ctx->topo : nullptr;
    P.xt = (const R*)ctx->d_xt;
    P.ainv = (const R*)ctx->d_ainv;
    P.push = nullptr;        // set by the marching launcher when the halo is fused into the stage
    P.partials = nullptr;    // ... when the conservation sums are fused into the stage
    P.ticket = nullptr;
    P.diag_out = nullptr;
    P.sums_in = 0;
    P.n = ctx->n, P.g = ctx->g, P.m = ctx->m, P.p = ctx->p, P.T = ctx->T;
    P.tile_ids = ctx->plain_order ? nullptr : ctx->d_tile_ids;
    P.t0 = ctx->t0;
    P.part = part;
    P.advect_only = ctx->cfg.advect_only;
    // a convex pair (a + b == 1) must stay one after rounding to R, or every step leaks a + b - 1 of the
    // mass (fp32: (float)(1/3) + (float)(2/3) = 1 + 6e-8); 1 - b is exact for b in [1/4, 1]
    P.b = (R)b, P.dt = (R)dt;
    P.a = (a != 0.0 && a + b > 1.0 - 1e-12 && a + b < 1.0 + 1e-12) ? R(1) - P.b : (R)a;
    P.rd = (R)(1.0 / ctx->delta);
    P.r2d = (R)(0.5 / ctx->delta);
    P.a2 = (R)(ctx->radius * ctx->radius);
    P.ia2 = (R)(1.0 / (ctx->radius * ctx->radius));
    P.grav = (R)ctx->gravity;
    P.two_omega = (R)(2.0 * ctx->omega);
    P.darea = (R)(ctx->delta * ctx->delta);
    for (int f = 0; f < 6; ++f)
        for (int k = 0; k < 3; ++k) P.cf[f][k] = (R)ctx->cf[f][k];
    return P;
}

// Work decomposition shared by the marching kernels: one warp per (tile, strip, chunk).
// Strip height trades the 4-row prologue of every strip against whole "waves" of warps: with W warp
// slots on the GPU the launch takes ceil(items / W) rounds of (rows + prologue) row-times, so the
// height is chosen to minimise that product (CS_MARCH_ROWS overrides).
template <typename Kernel>
MarchGeom march_geometry(cs_ctx* ctx, Kernel kernel, int V, int n, int g, int m, int T, int lo, int hi, size_t smem) {
    MarchGeom G;
    G.xlo = G.ylo = lo, G.xhi = G.yhi = hi;
    G.prefetch = ctx->tune.march_prefetch;
    G.ngroups = m / V;
    G.grp0 = (lo + g) / V;
    const int grp_last = (hi - 1 + g) / V;
    G.nchunks = (grp_last - G.grp0 + 1 + CS_MARCH_OUT_LANES - 1) / CS_MARCH_OUT_LANES;
    static std::map<std::tuple<int, const void*, size_t>, int> occupancy;  // per device, kernel instantiation and smem size
    static std::mutex occupancy_mutex;                                      // contexts may be driven from several host threads
    std::lock_guard<std::mutex> lock(occupancy_mutex);
    const auto key = std::make_tuple(ctx->device, reinterpret_cast<const void*>(kernel), smem);
    auto it = occupancy.find(key);
    if (it == occupancy.end()) {
        int bps = 1;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, kernel, CS_MARCH_THREADS, smem) != cudaSuccess || bps < 1) {
            cudaGetLastError();
            bps = 1;
        }
        it = occupancy.emplace(key, bps).first;
    }
    const int blocks_per_sm = it->second;
    // the inner part of a split stage (lo > 0) shares the GPU with the rim / pack / NCCL kernels of the side
    // stream: leave them a few SMs, otherwise the displaced warps form a second wave
    const int reserve = lo > 0 ? ctx->tune.reserve_sms : 0;
    const int sms = ctx->num_sms - reserve > 8 ? ctx->num_sms - reserve : ctx->num_sms;
    const long long slots = (long long)sms * blocks_per_sm * (CS_MARCH_THREADS / 32);
    const int span = hi - lo;
    int best_rows = span < 4 ? span : 4;
    double best_cost = 1e300;
    for (int rows = 4; rows <= 256 && rows <= span + 3; rows += 4) {
        const int r = rows < span ? rows : span;
        const long long items = (long long)G.nchunks * ((span + r - 1) / r) * T;
        const long long waves = (items + slots - 1) / slots;
        const double cost = (double)waves * (r + 5);
        if (cost < best_cost) best_cost = cost, best_rows = r;
    }
    G.rows = ctx->tune.march_rows > 0 ? ctx->tune.march_rows : best_rows;
    if (G.rows < 1) G.rows = 1;
    G.nstrips = (span + G.rows - 1) / G.rows;
    G.nitems = G.nchunks * G.nstrips * T;
    (void)n;
    return G;
}

}  // namespace cs_launch
