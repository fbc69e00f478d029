// K2 -- fused Runge-Kutta stage:  out = a q0 + b (qs + dt L(qs)).
#include "cs_internal.cuh"
#include "cs_cell.cuh"

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
    P.push = nullptr;
    P.partials = nullptr;
    P.n = ctx->n, P.g = ctx->g, P.m = ctx->m, P.p = ctx->p, P.T = ctx->T, P.t0 = ctx->t0;
    P.part = part;
    P.advect_only = ctx->cfg.advect_only;
    P.a = (R)a, P.b = (R)b, P.dt = (R)dt;
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

template <typename R, int LIM>
int launch_naive(cs_ctx* ctx, const StageParams<R>& P, cudaStream_t s) {
    dim3 block(32, 8);
    dim3 grid((P.n + 31) / 32, (P.n + 7) / 8, P.T);
    stage_naive_kernel<R, LIM><<<grid, block, 0, s>>>(P);
    ctx->launches++;
    CS_CUDA(ctx, cudaGetLastError());
    return CS_OK;
}

template <typename R>
int launch_typed(cs_ctx* ctx, const void* const q0[3], const void* const qs[3], void* const qout[3], double a, double b,
                 double dt, int part, double* diag_out, cudaStream_t s) {
    StageParams<R> P = make_params<R>(ctx, q0, qs, qout, a, b, dt, part);
    (void)diag_out;
    switch (ctx->cfg.limiter) {
        case CS_LIM_NONE: return launch_naive<R, CS_LIM_NONE>(ctx, P, s);
        case CS_LIM_MINMOD: return launch_naive<R, CS_LIM_MINMOD>(ctx, P, s);
        default: return launch_naive<R, CS_LIM_MC>(ctx, P, s);
    }
}

}  // namespace

int cs_launch_stage(cs_ctx* ctx, const void* const q0[3], const void* const qs[3], void* const qout[3], double a, double b,
                    double dt, int part, double* diag_out, cudaStream_t s) {
    if (ctx->dtype == CS_F64) return launch_typed<double>(ctx, q0, qs, qout, a, b, dt, part, diag_out, s);
    return launch_typed<float>(ctx, q0, qs, qout, a, b, dt, part, diag_out, s);
}
