// Context life cycle, validation, metric tables, error text.
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "cs_internal.cuh"

static thread_local std::string g_thread_error;

int cs_fail(const cs_ctx* ctx, int code, const char* fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_thread_error = buf;
    if (ctx) ctx->err = buf;
    return code;
}

void cs_drop_graphs(cs_ctx* ctx) {
    for (StepGraph& g : ctx->graphs)
        if (g.exec) {
            cudaGraphExecDestroy(g.exec);
            g = StepGraph();
        }
}

extern "C" int cs_abi_version(void) { return CS_ABI_VERSION; }

extern "C" const char* cs_last_error(const cs_ctx* ctx) { return ctx ? ctx->err.c_str() : g_thread_error.c_str(); }

extern "C" int64_t cs_launch_count(const cs_ctx* ctx) { return ctx ? ctx->launches : 0; }

extern "C" int cs_ctx_set_option(cs_ctx* ctx, const char* name, int value) {
    if (!ctx || !name) return cs_fail(ctx, CS_E_INVALID, "cs_ctx_set_option: NULL argument");
    struct { const char* key; int* slot; } opts[] = {
        {"march_prefetch", &ctx->tune.march_prefetch}, {"march_rows", &ctx->tune.march_rows}, {"march_vec", &ctx->tune.march_vec},
        {"march_unit", &ctx->tune.march_unit},         {"reserve_sms", &ctx->tune.reserve_sms}, {"tma_ring", &ctx->tune.tma_ring},
        {"peer_timeout_ms", &ctx->tune.peer_timeout_ms}, {"fused_halo", &ctx->tune.fused_halo},         {"step_graph", &ctx->tune.step_graph},   {"fused_diag", &ctx->tune.fused_diag},
        {"step_kernel", &ctx->tune.step_kernel},
    };
    for (auto& o : opts)
        if (std::strcmp(o.key, name) == 0) {
            *o.slot = value;
            cs_drop_graphs(ctx);                          // captured steps have the old choice baked in
            return CS_OK;
        }
    return cs_fail(ctx, CS_E_INVALID, "cs_ctx_set_option: unknown option '%s'", name);
}

extern "C" int cs_stage_kernel_info(const cs_ctx* ctx, int32_t* last_variant, int64_t* naive_fallbacks) {
    if (!ctx) return cs_fail(ctx, CS_E_INVALID, "cs_stage_kernel_info: NULL context");
    if (last_variant) *last_variant = ctx->last_stage_kernel;
    if (naive_fallbacks) *naive_fallbacks = ctx->naive_fallbacks;
    return CS_OK;
}

namespace {

template <typename R>
int upload_table(cs_ctx* ctx, void** dptr, const std::vector<double>& v) {
    std::vector<R> tmp(v.begin(), v.end());
    if (cudaMalloc(dptr, tmp.size() * sizeof(R)) != cudaSuccess)
        return cs_fail(ctx, CS_E_NOMEM, "cudaMalloc of a metric table failed");
    CS_CUDA(ctx, cudaMemcpy(*dptr, tmp.data(), tmp.size() * sizeof(R), cudaMemcpyHostToDevice));
    return CS_OK;
}

}  // namespace

// device copies of the tangent table ctx->xt (N + 2g entries) in the state's dtype and in fp64, and of 1 / (1 + X^2)
int cs_upload_metric(cs_ctx* ctx) {
    std::vector<double> ainv(ctx->xt.size());
    for (size_t I = 0; I < ctx->xt.size(); ++I) ainv[I] = 1.0 / (1.0 + ctx->xt[I] * ctx->xt[I]);
    for (void** q : {&ctx->d_xt, &ctx->d_ainv, (void**)&ctx->d_xt64})
        if (*q) cudaFree(*q), *q = nullptr;
    int rc;
    if ((rc = upload_table<double>(ctx, (void**)&ctx->d_xt64, ctx->xt))) return rc;
    if (ctx->dtype == CS_F64) {
        if ((rc = upload_table<double>(ctx, &ctx->d_xt, ctx->xt))) return rc;
        if ((rc = upload_table<double>(ctx, &ctx->d_ainv, ainv))) return rc;
    } else {
        if ((rc = upload_table<float>(ctx, &ctx->d_xt, ctx->xt))) return rc;
        if ((rc = upload_table<float>(ctx, &ctx->d_ainv, ainv))) return rc;
    }
    return CS_OK;
}

namespace {

int build_tables(cs_ctx* ctx, const double* tangents) {
    const int N = ctx->N, g = ctx->g;
    std::vector<double> x(N);
    if (tangents) {
        for (int i = 0; i < N; ++i) x[i] = tangents[i];
    } else {
        const double pi = 3.14159265358979323846;
        for (int i = 0; i < N; ++i) x[i] = std::tan(-0.25 * pi + (i + 0.5) * (0.5 * pi / N));
        for (int i = 0; i < N / 2; ++i) x[i] = -x[N - 1 - i];   // exactly antisymmetric
        if (N % 2) x[N / 2] = 0.0;
    }
    ctx->x_base = x;
    ctx->xt.resize(N + 2 * g);
    for (int I = 0; I < N + 2 * g; ++I) ctx->xt[I] = x[((I - g) % N + N) % N];      // periodic continuation
    int rc;
    if ((rc = cs_upload_metric(ctx))) return rc;
    static const double FRd[6][3][3] = {
        {{0, 0, 1}, {1, 0, 0}, {0, 1, 0}},   {{0, 1, 0}, {-1, 0, 0}, {0, 0, 1}},
        {{-1, 0, 0}, {0, -1, 0}, {0, 0, 1}}, {{0, -1, 0}, {1, 0, 0}, {0, 0, 1}},
        {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}},   {{0, 0, -1}, {-1, 0, 0}, {0, 1, 0}},
    };
    for (int f = 0; f < 6; ++f)
        for (int k = 0; k < 3; ++k)
            ctx->cf[f][k] = ctx->axis[0] * FRd[f][k][0] + ctx->axis[1] * FRd[f][k][1] + ctx->axis[2] * FRd[f][k][2];
    return CS_OK;
}

}  // namespace

extern "C" int cs_ctx_create(const cs_config* cfg, cs_ctx** out) {
    if (!cfg || !out) return cs_fail(nullptr, CS_E_INVALID, "cs_ctx_create: NULL argument");
    *out = nullptr;
    if (cfg->abi_version != CS_ABI_VERSION)
        return cs_fail(nullptr, CS_E_INVALID, "cs_config.abi_version = %d, library has %d", cfg->abi_version,
                       CS_ABI_VERSION);
    const int N = cfg->N, p = cfg->tiles_per_edge, g = cfg->halo;
    if (p < 1) return cs_fail(nullptr, CS_E_INVALID, "tiles_per_edge = %d must be >= 1", p);
    if (N < 2 || N % p) return cs_fail(nullptr, CS_E_INVALID, "N = %d must be >= 2 and divisible by tiles_per_edge = %d", N, p);
    const int n = N / p;
    if (g < 1 || g > CS_MAX_HALO || g > n)
        return cs_fail(nullptr, CS_E_INVALID, "halo = %d must satisfy 1 <= halo <= min(%d, n = %d)", g, CS_MAX_HALO, n);
    if (cfg->dtype != CS_F32 && cfg->dtype != CS_F64) return cs_fail(nullptr, CS_E_INVALID, "unknown dtype %d", cfg->dtype);
    if (cfg->rk < CS_RK_EULER || cfg->rk > CS_RK_SSPRK3) return cs_fail(nullptr, CS_E_INVALID, "unknown rk scheme %d", cfg->rk);
    if (cfg->limiter < CS_LIM_NONE || cfg->limiter > CS_LIM_MC) return cs_fail(nullptr, CS_E_INVALID, "unknown limiter %d", cfg->limiter);
    if (cfg->kernel < CS_KERNEL_AUTO || cfg->kernel > CS_KERNEL_TMA) return cs_fail(nullptr, CS_E_INVALID, "unknown kernel variant %d", cfg->kernel);
    const int T_all = 6 * p * p;
    const int T = cfg->tile_count == 0 ? T_all : cfg->tile_count;
    if (cfg->tile_begin < 0 || T < 1 || cfg->tile_begin + T > T_all)
        return cs_fail(nullptr, CS_E_INVALID, "tile range [%d, %d) is outside [0, %d)", cfg->tile_begin, cfg->tile_begin + T, T_all);
    const size_t esize = cfg->dtype == CS_F64 ? 8 : 4;
    const int m = n + 2 * g;
    if ((double)T * m * m >= 2147483647.0)
        return cs_fail(nullptr, CS_E_UNSUPPORTED, "shard of %d tiles of %d^2 cells exceeds 32-bit element offsets", T, m);

    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return cs_fail(nullptr, CS_E_CUDA, "no CUDA device available; this library has no CPU fallback");
    }
    if (cfg->device < 0 || cfg->device >= ndev)
        return cs_fail(nullptr, CS_E_INVALID, "device %d out of range [0, %d)", cfg->device, ndev);

    CS_ON_DEVICE(nullptr, cfg->device);              // the caller's current device is restored on every return path
    cs_ctx* ctx = new cs_ctx();
    ctx->cfg = *cfg;
    ctx->cfg.tangents = nullptr;
    ctx->N = N, ctx->p = p, ctx->n = n, ctx->g = g, ctx->m = m;
    ctx->T_all = T_all, ctx->t0 = cfg->tile_begin, ctx->T = T;
    ctx->order.resize(T_all), ctx->inv.assign(T_all, -1);
    for (int k = 0; k < T_all; ++k) ctx->order[k] = cfg->tile_order ? cfg->tile_order[k] : k;
    for (int k = 0; k < T_all; ++k) {
        const int t = ctx->order[k];
        if (t < 0 || t >= T_all || ctx->inv[t] >= 0) {
            delete ctx;
            return cs_fail(nullptr, CS_E_INVALID, "tile_order is not a permutation of 0 .. %d (entry %d = %d)", T_all - 1, k, t);
        }
        ctx->inv[t] = k;
    }
    ctx->cfg.tile_order = nullptr;
    for (int k = 0; k < T; ++k)
        if (ctx->order[cfg->tile_begin + k] != cfg->tile_begin + k) ctx->plain_order = false;
    ctx->dtype = cfg->dtype, ctx->esize = esize, ctx->device = cfg->device;
    ctx->field_elems = (size_t)T * m * m;
    ctx->pitch16 = ((size_t)m * esize) % 16 == 0;   // vectorised / TMA stage kernels need it
    ctx->radius = cfg->radius != 0 ? cfg->radius : 6.37122e6;
    ctx->omega = cfg->omega != 0 ? cfg->omega : 7.292e-5;
    ctx->gravity = cfg->gravity != 0 ? cfg->gravity : 9.80616;
    double an = std::sqrt(cfg->axis[0] * cfg->axis[0] + cfg->axis[1] * cfg->axis[1] + cfg->axis[2] * cfg->axis[2]);
    if (an == 0) ctx->axis[0] = 0, ctx->axis[1] = 0, ctx->axis[2] = 1;
    else for (int k = 0; k < 3; ++k) ctx->axis[k] = cfg->axis[k] / an;
    ctx->delta = 0.5 * 3.14159265358979323846 / N;

    auto bail = [&](int rc) {
        std::string msg = ctx->err;
        cs_ctx_destroy(ctx);
        g_thread_error = msg;
        return rc;
    };
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, ctx->device) != cudaSuccess) return bail(cs_fail(ctx, CS_E_CUDA, "cudaGetDeviceProperties failed"));
    if (prop.major != 10)
        return bail(cs_fail(ctx, CS_E_CUDA, "device %d is sm_%d%d; this library carries sm_100a code only", ctx->device, prop.major, prop.minor));
    ctx->num_sms = prop.multiProcessorCount;
    {
        int khz = 0;                                      // (cudaDeviceProp::clockRate is gone from newer toolkits)
        if (cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, ctx->device) == cudaSuccess && khz > 0) ctx->clock_khz = khz;
    }

    int rc;
    if ((rc = build_tables(ctx, cfg->tangents))) return bail(rc);
    if (cudaMalloc((void**)&ctx->d_tile_ids, (size_t)T * sizeof(int32_t)) != cudaSuccess)
        return bail(cs_fail(ctx, CS_E_NOMEM, "cudaMalloc of the tile table failed"));
    if (cudaMemcpy(ctx->d_tile_ids, ctx->order.data() + ctx->t0, (size_t)T * sizeof(int32_t), cudaMemcpyHostToDevice) != cudaSuccess)
        return bail(cs_fail(ctx, CS_E_CUDA, "upload of the tile table failed"));
    if ((rc = cs_build_halo(ctx))) return bail(rc);
    const size_t bytes = ctx->field_elems * esize;
    if (g >= 2) {
        const int nbuf = cfg->rk == CS_RK_EULER ? 1 : 2;
        for (int k = 0; k < 3; ++k) {
            if (cudaMalloc(&ctx->qa[k], bytes) != cudaSuccess) return bail(cs_fail(ctx, CS_E_NOMEM, "cudaMalloc of RK scratch (%zu bytes) failed", bytes));
            cudaMemsetAsync(ctx->qa[k], 0, bytes, 0);
            if (nbuf > 1) {
                if (cudaMalloc(&ctx->qb[k], bytes) != cudaSuccess) return bail(cs_fail(ctx, CS_E_NOMEM, "cudaMalloc of RK scratch (%zu bytes) failed", bytes));
                cudaMemsetAsync(ctx->qb[k], 0, bytes, 0);
            }
        }
    }
    if (cudaMalloc(&ctx->topo, bytes) != cudaSuccess) return bail(cs_fail(ctx, CS_E_NOMEM, "cudaMalloc of topography failed"));
    cudaMemsetAsync(ctx->topo, 0, bytes, 0);
    ctx->partial_cap = 65536;   // one (mass, energy) pair per block of the reduction kernel or per work item of a fused stage
    // three regions of partials and three tickets: [0] stage kernel (whole stage or inner part), [1] rim part (it may
    // run beside the inner part on another stream), [2] the stand-alone reduction kernel
    if (cudaMalloc((void**)&ctx->d_partials, 3 * (size_t)ctx->partial_cap * 2 * sizeof(double)) != cudaSuccess ||
        cudaMalloc((void**)&ctx->d_diag, 2 * sizeof(double)) != cudaSuccess ||
        cudaMalloc((void**)&ctx->d_ticket, 3 * sizeof(unsigned int)) != cudaSuccess)
        return bail(cs_fail(ctx, CS_E_NOMEM, "cudaMalloc of reduction scratch failed"));
    cudaMemsetAsync(ctx->d_ticket, 0, 3 * sizeof(unsigned int), 0);
    if (cudaDeviceSynchronize() != cudaSuccess) return bail(cs_fail(ctx, CS_E_CUDA, "context set-up failed: %s", cudaGetErrorString(cudaGetLastError())));
    if (const char* e = std::getenv("CS_PDL")) ctx->pdl = *e != '0';
    // the tuning environment is read here, once: no getenv on any launch path
    auto env_int = [](const char* name, int dflt) {
        const char* v = std::getenv(name);
        return v && *v ? std::atoi(v) : dflt;
    };
    CsTuning& tu = ctx->tune;
    tu.march_prefetch = env_int("CS_MARCH_PREFETCH", tu.march_prefetch);
    tu.march_rows = env_int("CS_MARCH_ROWS", tu.march_rows);
    tu.march_vec = env_int("CS_MARCH_VEC", tu.march_vec);
    tu.march_unit = env_int("CS_MARCH_UNIT", tu.march_unit);
    tu.reserve_sms = env_int("CS_MARCH_RESERVE_SMS", tu.reserve_sms);
    tu.tma_ring = env_int("CS_TMA_RING", tu.tma_ring);
    tu.fused_halo = env_int("CS_FUSED_HALO", tu.fused_halo);
    tu.peer_timeout_ms = env_int("CS_PEER_TIMEOUT_MS", tu.peer_timeout_ms);
    tu.step_graph = env_int("CS_STEP_GRAPH", tu.step_graph);
    tu.fused_diag = env_int("CS_FUSED_DIAG", tu.fused_diag);
    tu.step_kernel = env_int("CS_STEP_KERNEL", tu.step_kernel);
    tu.step_debug = env_int("CS_STEP_DEBUG", tu.step_debug);
    *out = ctx;
    return CS_OK;
}

extern "C" int cs_ctx_destroy(cs_ctx* ctx) {
    if (!ctx) return CS_OK;
    CsDeviceGuard guard(ctx->device);
    cudaDeviceSynchronize();
    cs_peer_destroy(ctx);
    cs_drop_graphs(ctx);
    if (ctx->cap_stream) cudaStreamDestroy(ctx->cap_stream);
    for (cudaStream_t s : {ctx->pipe.in, ctx->pipe.compute, ctx->pipe.out})
        if (s) cudaStreamDestroy(s);
    for (int k = 0; k < 2; ++k)
        for (cudaEvent_t e : {ctx->pipe.in_done[k], ctx->pipe.compute_done[k], ctx->pipe.out_done[k]})
            if (e) cudaEventDestroy(e);
    for (int k = 0; k < 3; ++k)
        if (ctx->pipe.state[1][k]) cudaFree(ctx->pipe.state[1][k]);
    if (ctx->pipe.diag) cudaFree(ctx->pipe.diag);
    if (ctx->pipe.diag_pinned) cudaFreeHost(ctx->pipe.diag_pinned);
    void* ptrs[] = {ctx->d_xt, ctx->d_ainv, ctx->d_xt64, ctx->d_halo_dst, ctx->d_halo_src, ctx->d_halo_code, ctx->d_rot, ctx->d_push, ctx->d_tile_ids, ctx->d_interp_dst, ctx->d_interp_src, ctx->d_interp_w, ctx->d_interp_rot,
                    ctx->d_send_src, ctx->d_recv_dst, ctx->d_recv_code, ctx->d_send_off, ctx->d_recv_off, ctx->qa[0], ctx->qa[1], ctx->qa[2],
                    ctx->qb[0], ctx->qb[1], ctx->qb[2], ctx->topo, ctx->host_state[0], ctx->host_state[1], ctx->host_state[2], ctx->d_partials, ctx->d_diag, ctx->d_ticket};
    for (void* q : ptrs)
        if (q) cudaFree(q);
    delete ctx;
    return CS_OK;
}
