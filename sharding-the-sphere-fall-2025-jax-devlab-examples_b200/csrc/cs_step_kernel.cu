// One launch per time step: step_march_kernel (cs_stage_march.cuh) behind cs_step.
//
// Taken when the context owns the whole sphere on one device (every neighbour tile is local), the layout can be
// vectorised, no rotation table is installed, the scheme has two or three stages, and the caller wants either no sums
// or the sums of the entry state.  Everything else goes through the one-stage-per-launch path.
#include "cs_internal.cuh"
#include "cs_cell.cuh"
#include "cs_stage_march.cuh"
#include "cs_stage_launch.cuh"

using namespace cs_launch;

namespace {

template <typename R, int V, int LIM, int MINB, bool UNIT1>
int launch_step(cs_ctx* ctx, StepParams<R>& S, double* diag_out, cudaStream_t s) {
    auto go = [&](auto kernel) -> int {
        const MarchGeom G = march_geometry(ctx, kernel, V, ctx->n, ctx->g, ctx->m, ctx->T, 0, ctx->n, 0);
        const int warps = CS_MARCH_THREADS / 32;
        int bps = 0;
        CS_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, kernel, CS_MARCH_THREADS, 0));
        const int resident = bps * ctx->num_sms;                       // a cooperative grid must be co-resident
        if (resident < 1) return cs_fail(ctx, CS_E_CUDA, "the step kernel does not fit on an SM");
        const int want = (G.nitems + warps - 1) / warps;
        const int blocks = want < resident ? want : resident;
        if ((int64_t)blocks * warps > ctx->partial_cap) return cs_fail(ctx, CS_E_UNSUPPORTED, "step kernel: too many warps for the partial sums");
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(blocks), cfg.blockDim = dim3(CS_MARCH_THREADS), cfg.dynamicSmemBytes = 0, cfg.stream = s;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeCooperative;
        attr[0].val.cooperative = 1;
        cfg.attrs = attr, cfg.numAttrs = 1;
        CS_CUDA(ctx, cudaLaunchKernelEx(&cfg, kernel, S, G));
        ctx->launches++;
        return CS_OK;
    };
    if (diag_out) {
        S.st[0].partials = ctx->d_partials, S.st[0].ticket = ctx->d_ticket, S.st[0].diag_out = diag_out, S.st[0].sums_in = 1;
        return go(step_march_kernel<R, V, LIM, MINB, UNIT1, 2>);
    }
    return go(step_march_kernel<R, V, LIM, MINB, UNIT1, 0>);
}

template <typename R, int LIM>
int launch_step_lim(cs_ctx* ctx, StepParams<R>& S, double* diag_out, cudaStream_t s) {
    if constexpr (sizeof(R) == 4) {
        if (ctx->tune.march_vec == 4 && ctx->m % 4 == 0) return launch_step<R, 4, LIM, 2, false>(ctx, S, diag_out, s);
        return launch_step<R, 2, LIM, 4, false>(ctx, S, diag_out, s);
    } else {
        return launch_step<R, 2, LIM, 2, true>(ctx, S, diag_out, s);
    }
}

template <typename R>
int launch_step_typed(cs_ctx* ctx, void* h, void* u, void* v, double dt, const double (*pairs)[2], int ns, double* diag_out,
                      cudaStream_t s) {
    StepParams<R> S;
    void* q0[3] = {h, u, v};
    const void* cur[3] = {h, u, v};
    for (int k = 0; k < ns; ++k) {
        void** dst = (k == ns - 1) ? q0 : (k % 2 == 0 ? ctx->qa : ctx->qb);
        S.st[k] = make_params<R>(ctx, q0, cur, dst, pairs[k][0], pairs[k][1], dt, 2);
        S.st[k].push = ctx->d_push;
        for (int c = 0; c < 3; ++c) cur[c] = dst[c];
    }
    for (int k = ns; k < 3; ++k) S.st[k] = S.st[ns - 1];
    S.nstages = ns;
    S.debug = ctx->tune.step_debug;
    S.trace = reinterpret_cast<unsigned long long*>(ctx->d_partials + 4 * (size_t)ctx->partial_cap);   // the reduction kernel's region
    S.halo_dst = ctx->d_halo_dst, S.halo_src = ctx->d_halo_src, S.halo_code = ctx->d_halo_code, S.halo_count = ctx->n_local;
    switch (ctx->cfg.limiter) {
        case CS_LIM_NONE: return launch_step_lim<R, CS_LIM_NONE>(ctx, S, diag_out, s);
        case CS_LIM_MINMOD: return launch_step_lim<R, CS_LIM_MINMOD>(ctx, S, diag_out, s);
        default: return launch_step_lim<R, CS_LIM_MC>(ctx, S, diag_out, s);
    }
}

}  // namespace

bool cs_step_kernel_eligible(const cs_ctx* ctx, const void* h, const void* u, const void* v, const double* diag_out) {
    if (!ctx->tune.step_kernel) return false;
    if (ctx->cfg.kernel != CS_KERNEL_AUTO && ctx->cfg.kernel != CS_KERNEL_MARCH) return false;
    if (ctx->cfg.rk != CS_RK_SSPRK2 && ctx->cfg.rk != CS_RK_SSPRK3) return false;
    if (ctx->T != ctx->T_all || !ctx->push_complete || !ctx->d_push || ctx->d_rot || ctx->d_interp_dst) return false;
    if ((int64_t)ctx->halo.size() != ctx->n_local) return false;
    if (diag_out && !ctx->cfg.sums_of_entry_state) return false;       // sums of the NEW state: the last-stage build
    if (diag_out && !ctx->tune.fused_diag) return false;
    if (ctx->dtype == CS_F64 && !ctx->tune.march_unit) return false;
    if (ctx->g < 2 || ctx->m % 2 || (ctx->field_elems / ctx->T * ctx->esize) % 16) return false;
    const void* ptrs[] = {h, u, v, ctx->qa[0], ctx->qa[1], ctx->qa[2], ctx->qb[0], ctx->qb[1], ctx->qb[2], ctx->topo};
    for (const void* q : ptrs)
        if (!aligned16(q)) return false;
    return true;
}

int cs_launch_step_kernel(cs_ctx* ctx, void* h, void* u, void* v, double dt, double* diag_out, cudaStream_t s) {
    double pairs[3][2] = {{0.0, 1.0}, {0.0, 0.0}, {0.0, 0.0}};
    int ns;
    if (ctx->cfg.rk == CS_RK_SSPRK2) {
        ns = 2, pairs[1][0] = 0.5, pairs[1][1] = 0.5;
    } else {
        ns = 3, pairs[1][0] = 0.75, pairs[1][1] = 0.25, pairs[2][0] = 1.0 - 2.0 / 3.0, pairs[2][1] = 2.0 / 3.0;
    }
    ctx->last_stage_kernel = CS_KERNEL_MARCH;
    if (ctx->dtype == CS_F64) return launch_step_typed<double>(ctx, h, u, v, dt, pairs, ns, diag_out, s);
    return launch_step_typed<float>(ctx, h, u, v, dt, pairs, ns, diag_out, s);
}

// Tuning aid (CS_STEP_DEBUG=16): per-block globaltimer stamps of the last step-kernel launch, [blocks][8].
extern "C" int cs_step_kernel_trace(cs_ctx* ctx, int max_blocks, unsigned long long* out_host) {
    if (!ctx || !out_host || max_blocks < 1) return cs_fail(ctx, CS_E_INVALID, "cs_step_kernel_trace: bad argument");
    if ((int64_t)max_blocks * 8 > 2 * (int64_t)ctx->partial_cap) return cs_fail(ctx, CS_E_INVALID, "cs_step_kernel_trace: too many blocks");
    CS_ON_DEVICE(ctx, ctx->device);
    CS_CUDA(ctx, cudaDeviceSynchronize());
    CS_CUDA(ctx, cudaMemcpy(out_host, ctx->d_partials + 4 * (size_t)ctx->partial_cap, (size_t)max_blocks * 8 * sizeof(unsigned long long),
                            cudaMemcpyDeviceToHost));
    return CS_OK;
}
