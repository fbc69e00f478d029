// K2 -- fused Runge-Kutta stage:  out = a q0 + b (qs + dt L(qs)).
#include "cs_internal.cuh"
#include "cs_cell.cuh"
#include "cs_reduce.cuh"
#include "cs_stage_march.cuh"
#include "cs_stage_tma.cuh"
#include "cs_stage_launch.cuh"

#include <cstdlib>
#include <map>
#include <mutex>
#include <tuple>
#include <utility>

using namespace cs_launch;

// resident blocks per SM the fp32 four-column build is compiled for (tuning aid: 3 caps the kernel at 168 registers)
#ifndef CS_MINB_F32V4
#define CS_MINB_F32V4 2
#endif
// ... and the same for its first-stage (UNIT) build, which needs fewer registers (no q0 in flight)
#ifndef CS_MINB_UNIT_F32V4
#define CS_MINB_UNIT_F32V4 CS_MINB_F32V4
#endif

namespace {

// ---------------------------------------------------------------------------------------------
// CS_KERNEL_NAIVE: one thread per interior cell, loads straight from global memory.
// ---------------------------------------------------------------------------------------------
template <typename R, int LIM>
__global__ void __launch_bounds__(256)
stage_naive_kernel(const __grid_constant__ StageParams<R> P) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y * blockDim.y + threadIdx.y;
    const int t = blockIdx.z;
    if (i >= P.n || j >= P.n) return;
    if (P.part != 2) {
        const bool inner = i >= P.g && i < P.n - P.g && j >= P.g && j < P.n - P.g;
        if (inner != (P.part == 0)) return;
    }
    cs_stage_cell<R, LIM>(P, t, j, i);
}


// Rim cells only (those within g of a tile edge): what peers' ghosts depend on.  Used for part 1
// of a split stage, so that the exchange can start while the inner cells are still being computed.
// DIAG: the launch also leaves the (mass, energy) sums over the rim cells in P.diag_out (all-fp64 per cell, one
// partial per block, the last block adds them in index order).
template <typename R, int LIM, bool DIAG>
__global__ void __launch_bounds__(256)
stage_rim_kernel(const __grid_constant__ StageParams<R> P) {
    const int n = P.n, g = P.g;
    const int per_tile = n * n - (n - 2 * g) * (n - 2 * g);
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    double mass = 0.0, energy = 0.0;
    if (idx < per_tile) {
        int j, i;
        if (idx < 2 * g * n) {            // bottom and top g rows, full width
            const int r = idx / n;
            i = idx - r * n;
            j = r < g ? r : n - 2 * g + r;
        } else {                          // left and right g columns of the remaining rows
            const int r = idx - 2 * g * n;
            const int jj = r / (2 * g), c = r - jj * (2 * g);
            j = g + jj;
            i = c < g ? c : n - 2 * g + c;
        }
        if constexpr (DIAG) {
            CellOut<R> o;
            cs_stage_cell<R, LIM>(P, blockIdx.y, j, i, &o);
            if (P.sums_in)
                cs_cell_sums((double)o.X, (double)o.Y, (double)P.a2, (double)P.grav, (double)o.h_in, (double)o.u_in, (double)o.v_in,
                             (double)o.hs, mass, energy);
            else
                cs_cell_sums((double)o.X, (double)o.Y, (double)P.a2, (double)P.grav, (double)o.h, (double)o.u, (double)o.v,
                             (double)o.hs, mass, energy);
        } else {
            cs_stage_cell<R, LIM>(P, blockIdx.y, j, i);
        }
    }
    if constexpr (DIAG) {
        const double w = (double)P.darea;
        mass *= w, energy *= w;
        cs_block_sum2<256>(mass, energy);
        cs_finish_partials<256>(P.partials, P.ticket, blockIdx.y * gridDim.x + blockIdx.x, gridDim.x * gridDim.y, mass, energy,
                                P.diag_out);
    }
}

template <typename R, int LIM>
int launch_naive(cs_ctx* ctx, const StageParams<R>& P, cudaStream_t s) {
    dim3 block(32, 8);
    dim3 grid((P.n + 31) / 32, (P.n + 7) / 8, P.T);
    stage_naive_kernel<R, LIM><<<grid, block, 0, s>>>(P);
    ctx->launches++;
    CS_CUDA(ctx, cudaGetLastError());
    return CS_OK;
}

template <typename R, int LIM>
int launch_rim(cs_ctx* ctx, const StageParams<R>& P, double* diag_out, int sums_in, cudaStream_t s) {
    const int per_tile = P.n * P.n - (P.n - 2 * P.g) * (P.n - 2 * P.g);
    dim3 grid((per_tile + 255) / 256, P.T);
    if (diag_out && (int64_t)grid.x * grid.y <= ctx->partial_cap) {
        // the rim may run beside the inner kernel of the same stage on another stream: its own partials and ticket
        StageParams<R> Q = P;
        Q.partials = ctx->d_partials + 2 * (size_t)ctx->partial_cap, Q.ticket = ctx->d_ticket + 1, Q.diag_out = diag_out;
        Q.sums_in = sums_in;
        stage_rim_kernel<R, LIM, true><<<grid, 256, 0, s>>>(Q);
        ctx->stage_diag_fused = true;
    } else {
        stage_rim_kernel<R, LIM, false><<<grid, 256, 0, s>>>(P);
    }
    ctx->launches++;
    CS_CUDA(ctx, cudaGetLastError());
    return CS_OK;
}

template <typename R, int V, int LIM, int MINB>
int launch_march(cs_ctx* ctx, const StageParams<R>& P, int lo, int hi, double* diag_out, int sums_in, cudaStream_t s) {
    const MarchGeom G = march_geometry(ctx, stage_march_kernel<R, V, LIM, MINB, false>, V, P.n, P.g, P.m, P.T, lo, hi, 0);
    const StageParams<R>& PP = P;
    const int warps = CS_MARCH_THREADS / 32;
    // First stage, (a, b) = (0, 1): a variant compiled without q0 and without the multiplications by b.  11 % fewer
    // instructions in both precisions.  fp64 gains either way (169 -> 154 us at C1536).  fp32 gains without the sums
    // (74.7 vs 80.9 us) and loses with the entry-state sums in the same launch (whole step 0.289 vs 0.286 ms): a
    // different schedule, not fewer stalls -- re-measured after each change of the kernel (tools/ab_stage.py, ab_step.py).
    // march_unit: 0 never; 1 (default) fp64 always, fp32 for the launches without sums; 2 always
    const bool with_sums = diag_out != nullptr && sums_in != 0;
    const bool unit_variant = sizeof(R) == 8 ? ctx->tune.march_unit >= 1
                                             : (ctx->tune.march_unit >= 2 || (ctx->tune.march_unit == 1 && !with_sums));
    constexpr int MU = (sizeof(R) == 4 && V == 4) ? CS_MINB_UNIT_F32V4 : MINB;
    if (MU != MINB && unit_variant && P.a == R(0) && P.b == R(1) && !P.push) {
        // the first-stage build is compiled for another occupancy: its own work decomposition
        const MarchGeom GU = march_geometry(ctx, stage_march_kernel<R, V, LIM, MU, false, true>, V, P.n, P.g, P.m, P.T, lo, hi, 0);
        const dim3 gridu((GU.nitems + warps - 1) / warps), blocku(CS_MARCH_THREADS);
        if (diag_out && sums_in && ctx->tune.fused_diag && GU.nitems <= ctx->partial_cap) {
            StageParams<R> Q = PP;
            Q.partials = ctx->d_partials, Q.ticket = ctx->d_ticket, Q.diag_out = diag_out, Q.sums_in = sums_in;
            CS_CUDA(ctx, cs_launch_dependent(stage_march_kernel<R, V, LIM, MU, false, true, 2>, gridu, blocku, 0, s, ctx->pdl, Q, GU));
            ctx->stage_diag_fused = true;
            ctx->launches++;
            return CS_OK;
        }
        if (!diag_out) {
            CS_CUDA(ctx, cs_launch_dependent(stage_march_kernel<R, V, LIM, MU, false, true>, gridu, blocku, 0, s, ctx->pdl, PP, GU));
            ctx->launches++;
            return CS_OK;
        }
    }
    const dim3 grid((G.nitems + warps - 1) / warps), block(CS_MARCH_THREADS);
    if (diag_out && !P.push && ctx->tune.fused_diag && G.nitems <= ctx->partial_cap) {
        // last stage of a step: the conservation sums of the output come out of the same launch (K3 fused into K2)
        // ... or, for a first stage, those of its input (the state the step starts from), which cost next to nothing
        StageParams<R> Q = PP;
        Q.partials = ctx->d_partials, Q.ticket = ctx->d_ticket, Q.diag_out = diag_out, Q.sums_in = sums_in;
        if (!sums_in) {
            CS_CUDA(ctx, cs_launch_dependent(stage_march_kernel<R, V, LIM, MINB, false, false, 1>, grid, block, 0, s, ctx->pdl, Q, G));
        } else {
            bool launched = false;
            if (unit_variant && P.a == R(0) && P.b == R(1)) {
                CS_CUDA(ctx, cs_launch_dependent(stage_march_kernel<R, V, LIM, MINB, false, true, 2>, grid, block, 0, s, ctx->pdl, Q, G));
                launched = true;
            }
            if (!launched)
                CS_CUDA(ctx, cs_launch_dependent(stage_march_kernel<R, V, LIM, MINB, false, false, 2>, grid, block, 0, s, ctx->pdl, Q, G));
        }
        ctx->stage_diag_fused = true;
    } else if (P.push) {
        CS_CUDA(ctx, cs_launch_dependent(stage_march_kernel<R, V, LIM, MINB, true>, grid, block, 0, s, ctx->pdl, P, G));
    } else {
        bool launched = false;
        if (unit_variant && P.a == R(0) && P.b == R(1)) {
            CS_CUDA(ctx, cs_launch_dependent(stage_march_kernel<R, V, LIM, MINB, false, true>, grid, block, 0, s, ctx->pdl, PP, G));
            launched = true;
        }
        if (!launched)
            CS_CUDA(ctx, cs_launch_dependent(stage_march_kernel<R, V, LIM, MINB, false>, grid, block, 0, s, ctx->pdl, PP, G));
    }
    ctx->launches++;
    CS_CUDA(ctx, cudaGetLastError());
    return CS_OK;
}

template <typename R, int V, int LIM, int MINB>
int launch_tma(cs_ctx* ctx, const StageParams<R>& P, int lo, int hi, cudaStream_t s) {
    const int warps = CS_MARCH_THREADS / 32;
    int ring = ctx->tune.tma_ring;
    ring = ring < 4 ? 4 : ring > CS_TMA_MAX_RING ? CS_TMA_MAX_RING : ring;
    const int narr = 3 + (P.a != R(0) ? 3 : 0) + (P.topo ? 1 : 0);
    const size_t smem = (size_t)warps * ((size_t)ring * narr * cs_tma_astride<R, V>() + CS_TMA_MAX_RING * 8);
    // the attribute is per device (and per instantiation): set it on every launch, it costs about a microsecond
    CS_CUDA(ctx, cudaFuncSetAttribute(stage_tma_kernel<R, V, LIM, MINB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const MarchGeom G = march_geometry(ctx, stage_tma_kernel<R, V, LIM, MINB>, V, P.n, P.g, P.m, P.T, lo, hi, smem);
    stage_tma_kernel<R, V, LIM, MINB><<<(G.nitems + warps - 1) / warps, CS_MARCH_THREADS, smem, s>>>(P, G, ring);
    ctx->launches++;
    CS_CUDA(ctx, cudaGetLastError());
    return CS_OK;
}

template <typename R, int LIM>
int launch_lim(cs_ctx* ctx, const StageParams<R>& P0, int part, double* diag_out, int sums_in, cudaStream_t s) {
    const int n = P0.n, g = P0.g;
    ctx->stage_pushed = false;
    ctx->stage_diag_fused = false;
    StageParams<R> P = P0;
    if (n <= 2 * g && part != 2) {
        // a tile this small has no inner cells: every cell is within g of an edge, so the rim is the whole tile
        if (part == 0) return CS_OK;
        part = 2, P.part = 2;
    }
    if (part == 1) return launch_rim<R, LIM>(ctx, P, diag_out, sums_in, s);
    const int lo = part == 0 ? g : 0, hi = part == 0 ? n - g : n;
    if (hi <= lo) return CS_OK;
    int variant = ctx->cfg.kernel;
    bool ok16 = aligned16(P.hq) && aligned16(P.uq) && aligned16(P.vq) && aligned16(P.ho) && aligned16(P.uo) && aligned16(P.vo) &&
                aligned16(P.h0) && aligned16(P.u0) && aligned16(P.v0) && aligned16(P.topo);
    const bool can_march = ok16 && g >= 2 && (P.m % 2) == 0 && ((size_t)P.m * P.m * sizeof(R)) % 16 == 0;
    const bool can_tma = can_march && ((size_t)P.m * sizeof(R)) % 16 == 0;
    if (variant == CS_KERNEL_AUTO) {
        variant = can_march ? CS_KERNEL_MARCH : CS_KERNEL_NAIVE;
        if (!can_march) ctx->naive_fallbacks++;       // visible through cs_stage_kernel_info: a 4x slower path
    }
    ctx->last_stage_kernel = variant;
    if (variant == CS_KERNEL_MARCH && !can_march)
        return cs_fail(ctx, CS_E_INVALID, "the marching stage kernel needs 16-byte aligned fields, halo >= 2 and an even row length");
    if (variant == CS_KERNEL_TMA && !can_tma)
        return cs_fail(ctx, CS_E_INVALID, "the TMA stage kernel needs 16-byte aligned fields and rows, and halo >= 2");
    if (variant == CS_KERNEL_NAIVE) return launch_naive<R, LIM>(ctx, P, s);
    if (variant == CS_KERNEL_TMA) {
        if constexpr (sizeof(R) == 4) {
            if (ctx->tune.march_vec != 2 && P.m % 4 == 0) return launch_tma<R, 4, LIM, 2>(ctx, P, lo, hi, s);
            return launch_tma<R, 2, LIM, 4>(ctx, P, lo, hi, s);
        } else {
            return launch_tma<R, 2, LIM, 2>(ctx, P, lo, hi, s);
        }
    }
    StageParams<R> Q = P;
    ctx->stage_pushed = false;
    if (part == 2 && ctx->cfg.fused_halo && ctx->push_complete && ctx->d_push && !ctx->d_rot && !ctx->d_interp_dst && ctx->tune.fused_halo) {
        Q.push = ctx->d_push;
        ctx->stage_pushed = true;   // the caller can skip the ghost-fill launch (corner blocks are not refreshed)
    }
    // fp32: four columns per thread (one 16-byte access per row and variable) when rows allow it, else two;
    // fp64: two.  CS_MARCH_VEC=2 forces the two-column fp32 variant (tuning aid).
    if constexpr (sizeof(R) == 4) {
        if (ctx->tune.march_vec == 4 && P.m % 4 == 0) return launch_march<R, 4, LIM, CS_MINB_F32V4>(ctx, Q, lo, hi, diag_out, sums_in, s);
        return launch_march<R, 2, LIM, 4>(ctx, Q, lo, hi, diag_out, sums_in, s);
    } else {
        return launch_march<R, 2, LIM, 2>(ctx, Q, lo, hi, diag_out, sums_in, s);
    }
}

template <typename R>
int launch_typed(cs_ctx* ctx, const void* const q0[3], const void* const qs[3], void* const qout[3], double a, double b,
                 double dt, int part, double* diag_out, int sums_in, cudaStream_t s) {
    StageParams<R> P = make_params<R>(ctx, q0, qs, qout, a, b, dt, part);
    switch (ctx->cfg.limiter) {
        case CS_LIM_NONE: return launch_lim<R, CS_LIM_NONE>(ctx, P, part, diag_out, sums_in, s);
        case CS_LIM_MINMOD: return launch_lim<R, CS_LIM_MINMOD>(ctx, P, part, diag_out, sums_in, s);
        default: return launch_lim<R, CS_LIM_MC>(ctx, P, part, diag_out, sums_in, s);
    }
}

}  // namespace

int cs_launch_stage(cs_ctx* ctx, const void* const q0[3], const void* const qs[3], void* const qout[3], double a, double b,
                    double dt, int part, double* diag_out, int sums_in, cudaStream_t s) {
    if (ctx->dtype == CS_F64) return launch_typed<double>(ctx, q0, qs, qout, a, b, dt, part, diag_out, sums_in, s);
    return launch_typed<float>(ctx, q0, qs, qout, a, b, dt, part, diag_out, sums_in, s);
}
