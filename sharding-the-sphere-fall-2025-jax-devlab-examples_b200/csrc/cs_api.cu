// Entry points that compose the kernels: topography, one RK stage, one step, host-buffer step.
#include <cstdlib>

#include <cstring>

#include "cs_internal.cuh"

namespace {

struct RkPair { double a, b; };

// Shu-Osher pairs; a is 1 - b so that a + b == 1 exactly in binary floating point
// (1/3 + 2/3 rounds to 1 - 2^-54 and would leak mass every step).
int rk_pairs(int rk, RkPair out[3]) {
    out[0] = {0.0, 1.0};
    if (rk == CS_RK_EULER) return 1;
    if (rk == CS_RK_SSPRK2) {
        out[1] = {0.5, 0.5};
        return 2;
    }
    out[1] = {0.75, 0.25};
    out[2] = {1.0 - 2.0 / 3.0, 2.0 / 3.0};
    return 3;
}

int check_step_ready(cs_ctx* ctx, const char* who) {
    if (ctx->g < 2) return cs_fail(ctx, CS_E_INVALID, "%s: the stencil needs halo >= 2, the context has %d", who, ctx->g);
    return CS_OK;
}

bool single_device(const cs_ctx* ctx) { return ctx->T == ctx->T_all; }

}  // namespace

extern "C" int cs_set_topography(cs_ctx* ctx, const void* hs, void* stream) {
    if (!ctx) return cs_fail(ctx, CS_E_INVALID, "cs_set_topography: NULL context");
    CS_ON_DEVICE(ctx, ctx->device);
    cudaStream_t s = (cudaStream_t)stream;
    cs_drop_graphs(ctx);
    if (!hs) {
        ctx->has_topo = false;
        return CS_OK;
    }
    CS_CUDA(ctx, cudaMemcpyAsync(ctx->topo, hs, ctx->field_elems * ctx->esize, cudaMemcpyDeviceToDevice, s));
    ctx->has_topo = true;
    // local ghosts now; remote ghosts (shards) are the caller's job via cs_halo_pack(nfields = 1)
    return cs_launch_halo(ctx, ctx->topo, nullptr, nullptr, 1, s);
}

namespace {
int check_stage_args(cs_ctx* ctx, const void* const q0[3], const void* const qs[3], void* const qout[3], double a, int part) {
    if (!ctx || !q0 || !qs || !qout) return cs_fail(ctx, CS_E_INVALID, "cs_rk_stage: NULL argument");
    int rc = check_step_ready(ctx, "cs_rk_stage");
    if (rc) return rc;
    if (part < 0 || part > 2) return cs_fail(ctx, CS_E_INVALID, "cs_rk_stage_part: part = %d must be 0, 1 or 2", part);
    for (int k = 0; k < 3; ++k) {
        if (!qs[k] || !qout[k] || (a != 0.0 && !q0[k])) return cs_fail(ctx, CS_E_INVALID, "cs_rk_stage: NULL field pointer");
        for (int l = 0; l < 3; ++l)
            if (qout[k] == qs[l]) return cs_fail(ctx, CS_E_INVALID, "cs_rk_stage: qout must not alias qs (stencil input)");
    }
    return CS_OK;
}
}  // namespace

extern "C" int cs_rk_stage_part_sums(cs_ctx* ctx, const void* const q0[3], const void* const qs[3], void* const qout[3], double a,
                                     double b, double dt, int part, int which, double* sums_out, void* stream) {
    int rc = check_stage_args(ctx, q0, qs, qout, a, part);
    if (rc) return rc;
    if (which != CS_SUMS_OF_OUTPUT && which != CS_SUMS_OF_INPUT)
        return cs_fail(ctx, CS_E_INVALID, "cs_rk_stage_part_sums: which = %d must be CS_SUMS_OF_OUTPUT or CS_SUMS_OF_INPUT", which);
    CS_ON_DEVICE(ctx, ctx->device);
    cudaStream_t s = (cudaStream_t)stream;
    if ((rc = cs_launch_stage(ctx, q0, qs, qout, a, b, dt, part, sums_out, which == CS_SUMS_OF_INPUT, s))) return rc;
    if (sums_out && !ctx->stage_diag_fused) {
        if (part == 0 && ctx->n <= 2 * ctx->g) {          // no inner cells at all: their sums are zero
            CS_CUDA(ctx, cudaMemsetAsync(sums_out, 0, 2 * sizeof(double), s));
            return CS_OK;
        }
        return cs_fail(ctx, CS_E_UNSUPPORTED,
                       "cs_rk_stage_part_sums: the selected stage kernel cannot sum over part %d of a stage (the marching kernel "
                       "and the rim kernel can); the stage itself has been enqueued", part);
    }
    return CS_OK;
}

extern "C" int cs_rk_stage_part(cs_ctx* ctx, const void* const q0[3], const void* const qs[3], void* const qout[3], double a,
                                double b, double dt, int part, void* stream) {
    return cs_rk_stage_part_sums(ctx, q0, qs, qout, a, b, dt, part, CS_SUMS_OF_OUTPUT, nullptr, stream);
}

extern "C" int cs_rk_stage(cs_ctx* ctx, const void* const q0[3], const void* const qs[3], void* const qout[3], double a,
                           double b, double dt, double* diag_out, void* stream) {
    int rc = check_stage_args(ctx, q0, qs, qout, a, 2);
    if (rc) return rc;
    CS_ON_DEVICE(ctx, ctx->device);
    cudaStream_t s = (cudaStream_t)stream;
    // the marching kernel produces the conservation sums of qout in the same launch; other variants leave them
    // to the reduction kernel
    if ((rc = cs_launch_stage(ctx, q0, qs, qout, a, b, dt, 2, diag_out, 0, s))) return rc;
    const bool fused = ctx->stage_diag_fused;
    if (!ctx->stage_pushed && (rc = cs_launch_halo(ctx, qout[0], qout[1], qout[2], 3, s))) return rc;
    if (diag_out && !fused) return cs_launch_diag(ctx, qout[0], qout[1], qout[2], diag_out, s);
    return CS_OK;
}

namespace {

// The launches of one step, in order, on stream s.
int enqueue_step(cs_ctx* ctx, void* h, void* u, void* v, double dt, double* diag_out, cudaStream_t s) {
    int rc;
    RkPair pr[3];
    const int ns = rk_pairs(ctx->cfg.rk, pr);
    void* q0[3] = {h, u, v};
    const void* cur[3] = {h, u, v};
    bool diag_fused = false;
    // where the conservation sums come from: the last stage (sums of the new state) or, when the context was created
    // with sums_of_entry_state, the first stage (sums of the state the step starts from: next to free)
    const bool entry = ctx->cfg.sums_of_entry_state != 0;
    const int ksum = entry ? 0 : ns - 1;
    for (int k = 0; k < ns; ++k) {
        // the last stage writes the caller's arrays in place (it reads them point-wise only);
        // a single-stage scheme has nothing else to read from, so it goes through scratch
        void** dst = (k == ns - 1 && ns > 1) ? q0 : (k % 2 == 0 ? ctx->qa : ctx->qb);
        // the last stage also produces the conservation sums of the new state when its kernel can (K3 fused into K2)
        rc = cs_launch_stage(ctx, q0, cur, dst, pr[k].a, pr[k].b, dt, 2, k == ksum ? diag_out : nullptr, entry, s);
        if (rc) return rc;
        if (k == ksum) diag_fused = ctx->stage_diag_fused;
        if (k == ksum && diag_out && !diag_fused && entry) {
            // a stage kernel that cannot sum: the reduction kernel on the entry state, before anything overwrites it
            if ((rc = cs_launch_diag(ctx, h, u, v, diag_out, s))) return rc;
            diag_fused = true;
        }
        if (!ctx->stage_pushed && (rc = cs_launch_halo(ctx, dst[0], dst[1], dst[2], 3, s))) return rc;
        for (int c = 0; c < 3; ++c) cur[c] = dst[c];
    }
    if (ns == 1)
        for (int c = 0; c < 3; ++c)
            CS_CUDA(ctx, cudaMemcpyAsync(q0[c], ctx->qa[c], ctx->field_elems * ctx->esize, cudaMemcpyDeviceToDevice, s));
    if (diag_out && !diag_fused) return cs_launch_diag(ctx, h, u, v, diag_out, s);
    return CS_OK;
}

// A step is 6-7 short launches; replaying them as one CUDA graph removes the gaps between them.
// Graphs are keyed by the arguments that are baked into the kernels' parameters.
StepGraph* find_or_capture(cs_ctx* ctx, void* h, void* u, void* v, double dt, double* diag_out) {
    for (StepGraph& g : ctx->graphs)
        if (g.exec && g.h == h && g.u == u && g.v == v && g.dt == dt && g.diag == diag_out && g.topo == ctx->has_topo) {
            g.stamp = ++ctx->graph_clock;
            return &g;
        }
    if (!ctx->cap_stream && cudaStreamCreateWithFlags(&ctx->cap_stream, cudaStreamNonBlocking) != cudaSuccess) {
        cudaGetLastError();
        return nullptr;
    }
    StepGraph* slot = nullptr;                            // an empty slot, else the least recently used one
    for (StepGraph& g : ctx->graphs)
        if (!g.exec) {
            slot = &g;
            break;
        }
    if (!slot) {
        slot = &ctx->graphs[0];
        for (StepGraph& g : ctx->graphs)
            if (g.stamp < slot->stamp) slot = &g;
    }
    if (slot->exec) cudaGraphExecDestroy(slot->exec), slot->exec = nullptr;
    const int64_t l0 = ctx->launches;
    if (cudaStreamBeginCapture(ctx->cap_stream, cudaStreamCaptureModeThreadLocal) != cudaSuccess) {
        cudaGetLastError();
        return nullptr;
    }
    const int rc = enqueue_step(ctx, h, u, v, dt, diag_out, ctx->cap_stream);
    cudaGraph_t graph = nullptr;
    const cudaError_t e = cudaStreamEndCapture(ctx->cap_stream, &graph);
    const int64_t n = ctx->launches - l0;
    ctx->launches = l0;                                   // capturing is not launching
    if (rc != CS_OK || e != cudaSuccess || !graph) {
        if (graph) cudaGraphDestroy(graph);
        cudaGetLastError();
        return nullptr;
    }
    const cudaError_t ei = cudaGraphInstantiate(&slot->exec, graph, 0);
    cudaGraphDestroy(graph);
    if (ei != cudaSuccess) {
        cudaGetLastError();
        slot->exec = nullptr;
        return nullptr;
    }
    slot->h = h, slot->u = u, slot->v = v, slot->dt = dt, slot->diag = diag_out, slot->topo = ctx->has_topo;
    slot->launches = n;
    slot->stamp = ++ctx->graph_clock;
    return slot;
}

}  // namespace

extern "C" int cs_step(cs_ctx* ctx, void* h, void* u, void* v, double dt, double* diag_out, void* stream) {
    if (!ctx || !h || !u || !v) return cs_fail(ctx, CS_E_INVALID, "cs_step: NULL argument");
    int rc = check_step_ready(ctx, "cs_step");
    if (rc) return rc;
    if (!single_device(ctx))
        return cs_fail(ctx, CS_E_INVALID,
                       "cs_step: the context owns %d of %d tiles; drive shards with cs_rk_stage_part + cs_halo_pack/unpack",
                       ctx->T, ctx->T_all);
    CS_ON_DEVICE(ctx, ctx->device);
    cudaStream_t s = (cudaStream_t)stream;
    // graph replay unless the caller is itself capturing this stream (no nested capture) or opted out
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(s, &cap) != cudaSuccess) cudaGetLastError(), cap = cudaStreamCaptureStatusNone;
    // a whole-sphere context steps in ONE launch when it can (no graph needed: there is nothing to stitch together)
    if (cs_step_kernel_eligible(ctx, h, u, v, diag_out)) return cs_launch_step_kernel(ctx, h, u, v, dt, diag_out, s);
    if (ctx->tune.step_graph && cap == cudaStreamCaptureStatusNone) {
        StepGraph* g = find_or_capture(ctx, h, u, v, dt, diag_out);
        if (!g && ctx->pdl) {                             // a driver that cannot capture programmatic edges: plain edges
            ctx->pdl = false;
            g = find_or_capture(ctx, h, u, v, dt, diag_out);
        }
        if (g) {
            CS_CUDA(ctx, cudaGraphLaunch(g->exec, s));
            ctx->launches += g->launches;
            return CS_OK;
        }
    }
    return enqueue_step(ctx, h, u, v, dt, diag_out, s);
}

extern "C" int cs_step_host(cs_ctx* ctx, void* h_host, void* u_host, void* v_host, double dt, int nsteps, double* diag_host,
                            int copy_back) {
    if (!ctx || !h_host || !u_host || !v_host) return cs_fail(ctx, CS_E_INVALID, "cs_step_host: NULL argument");
    if (nsteps < 0) return cs_fail(ctx, CS_E_INVALID, "cs_step_host: nsteps = %d must be >= 0", nsteps);
    int rc = check_step_ready(ctx, "cs_step_host");
    if (rc) return rc;
    if (!single_device(ctx)) return cs_fail(ctx, CS_E_INVALID, "cs_step_host needs a context that owns all tiles");
    CS_ON_DEVICE(ctx, ctx->device);
    const size_t bytes = ctx->field_elems * ctx->esize;
    // device copy of the caller's host state, allocated on first use and kept by the context
    void** d = ctx->host_state;
    for (int k = 0; k < 3; ++k)
        if (!d[k] && cudaMalloc(&d[k], bytes) != cudaSuccess)
            return cs_fail(ctx, CS_E_NOMEM, "cs_step_host: cudaMalloc of %zu bytes failed", bytes);
    cudaStream_t s = 0;
    void* hostp[3] = {h_host, u_host, v_host};
    auto cleanup = [&](int code) {
        cudaStreamSynchronize(s);
        return code;
    };
    for (int k = 0; k < 3; ++k)
        if (cudaMemcpyAsync(d[k], hostp[k], bytes, cudaMemcpyHostToDevice, s) != cudaSuccess)
            return cleanup(cs_fail(ctx, CS_E_CUDA, "cs_step_host: host-to-device copy failed"));
    if ((rc = cs_launch_halo(ctx, d[0], d[1], d[2], 3, s))) return cleanup(rc);
    for (int it = 0; it < nsteps; ++it)
        if ((rc = cs_step(ctx, d[0], d[1], d[2], dt, nullptr, s))) return cleanup(rc);
    if (diag_host) {
        if ((rc = cs_launch_diag(ctx, d[0], d[1], d[2], ctx->d_diag, s))) return cleanup(rc);
        if (cudaMemcpyAsync(diag_host, ctx->d_diag, 2 * sizeof(double), cudaMemcpyDeviceToHost, s) != cudaSuccess)
            return cleanup(cs_fail(ctx, CS_E_CUDA, "cs_step_host: device-to-host copy failed"));
    }
    if (copy_back)
        for (int k = 0; k < 3; ++k)
            if (cudaMemcpyAsync(hostp[k], d[k], bytes, cudaMemcpyDeviceToHost, s) != cudaSuccess)
                return cleanup(cs_fail(ctx, CS_E_CUDA, "cs_step_host: device-to-host copy failed"));
    cudaError_t e = cudaStreamSynchronize(s);
    if (e != cudaSuccess) return cs_fail(ctx, CS_E_CUDA, "cs_step_host: %s", cudaGetErrorString(e));
    return CS_OK;
}

namespace {

// Streams, events and the second device slot of cs_step_host_batch, created on first use.
int pipe_setup(cs_ctx* ctx) {
    cs_ctx::HostPipe& p = ctx->pipe;
    const size_t bytes = ctx->field_elems * ctx->esize;
    for (int k = 0; k < 3; ++k) {
        if (!ctx->host_state[k] && cudaMalloc(&ctx->host_state[k], bytes) != cudaSuccess)
            return cs_fail(ctx, CS_E_NOMEM, "cs_step_host_batch: cudaMalloc of %zu bytes failed", bytes);
        p.state[0][k] = ctx->host_state[k];
        if (!p.state[1][k] && cudaMalloc(&p.state[1][k], bytes) != cudaSuccess)
            return cs_fail(ctx, CS_E_NOMEM, "cs_step_host_batch: cudaMalloc of %zu bytes failed", bytes);
    }
    if (p.ready) return CS_OK;
    if (cudaMalloc((void**)&p.diag, 4 * sizeof(double)) != cudaSuccess)
        return cs_fail(ctx, CS_E_NOMEM, "cs_step_host_batch: cudaMalloc failed");
    for (cudaStream_t* s : {&p.in, &p.compute, &p.out})
        CS_CUDA(ctx, cudaStreamCreateWithFlags(s, cudaStreamNonBlocking));
    for (int k = 0; k < 2; ++k)
        for (cudaEvent_t* e : {&p.in_done[k], &p.compute_done[k], &p.out_done[k]})
            CS_CUDA(ctx, cudaEventCreateWithFlags(e, cudaEventDisableTiming));
    p.ready = true;
    return CS_OK;
}

}  // namespace

extern "C" int cs_step_host_batch(cs_ctx* ctx, int members, void* const* h_host, void* const* u_host, void* const* v_host,
                                  double dt, int nsteps, double* diag_host, int copy_back) {
    if (!ctx || !h_host || !u_host || !v_host) return cs_fail(ctx, CS_E_INVALID, "cs_step_host_batch: NULL argument");
    if (members < 0 || nsteps < 0)
        return cs_fail(ctx, CS_E_INVALID, "cs_step_host_batch: members = %d and nsteps = %d must be >= 0", members, nsteps);
    for (int i = 0; i < members; ++i)
        if (!h_host[i] || !u_host[i] || !v_host[i]) return cs_fail(ctx, CS_E_INVALID, "cs_step_host_batch: member %d has a NULL array", i);
    int rc = check_step_ready(ctx, "cs_step_host_batch");
    if (rc) return rc;
    if (!single_device(ctx)) return cs_fail(ctx, CS_E_INVALID, "cs_step_host_batch needs a context that owns all tiles");
    CS_ON_DEVICE(ctx, ctx->device);
    CS_CUDA(ctx, cudaDeviceSynchronize());                // earlier work on other streams may still use the slots
    if ((rc = pipe_setup(ctx))) return rc;
    cs_ctx::HostPipe& p = ctx->pipe;
    const size_t bytes = ctx->field_elems * ctx->esize;
    // A device-to-host copy into pageable memory blocks the calling thread until it is done, which would keep the
    // next member's upload from even being enqueued: the diagnostics go through a page-locked staging array.
    if (diag_host && members > p.diag_cap) {
        if (p.diag_pinned) cudaFreeHost(p.diag_pinned);
        p.diag_pinned = nullptr, p.diag_cap = 0;
        const int cap = members < 1024 ? 1024 : members;   // page-locking is slow: do it once for any ordinary batch
        if (cudaHostAlloc((void**)&p.diag_pinned, (size_t)cap * 2 * sizeof(double), cudaHostAllocDefault) != cudaSuccess)
            return cs_fail(ctx, CS_E_NOMEM, "cs_step_host_batch: cudaHostAlloc failed");
        p.diag_cap = cap;
    }
    auto drain = [&](int code) {
        cudaStreamSynchronize(p.in), cudaStreamSynchronize(p.compute), cudaStreamSynchronize(p.out);
        return code;
    };
#define CS_PIPE(call)                                                                                       \
    do {                                                                                                    \
        cudaError_t e_ = (call);                                                                            \
        if (e_ != cudaSuccess) return drain(cs_fail(ctx, CS_E_CUDA, "cs_step_host_batch: %s", cudaGetErrorString(e_))); \
    } while (0)
    for (int i = 0; i < members; ++i) {
        const int k = i & 1;
        void* const hostp[3] = {h_host[i], u_host[i], v_host[i]};
        void** d = p.state[k];
        // upload: the slot is free once member i - 2 has been stepped (and, with copy_back, downloaded)
        if (i >= 2) CS_PIPE(cudaStreamWaitEvent(p.in, copy_back ? p.out_done[k] : p.compute_done[k], 0));
        for (int c = 0; c < 3; ++c) CS_PIPE(cudaMemcpyAsync(d[c], hostp[c], bytes, cudaMemcpyHostToDevice, p.in));
        CS_PIPE(cudaEventRecord(p.in_done[k], p.in));
        // steps (members share the context's scratch, so they run one after the other on one stream)
        CS_PIPE(cudaStreamWaitEvent(p.compute, p.in_done[k], 0));
        if (i >= 2) CS_PIPE(cudaStreamWaitEvent(p.compute, p.out_done[k], 0));      // its diagnostics slot has been read
        if ((rc = cs_launch_halo(ctx, d[0], d[1], d[2], 3, p.compute))) return drain(rc);
        for (int it = 0; it < nsteps; ++it)
            if ((rc = cs_step(ctx, d[0], d[1], d[2], dt, nullptr, p.compute))) return drain(rc);
        if (diag_host && (rc = cs_launch_diag(ctx, d[0], d[1], d[2], p.diag + 2 * k, p.compute))) return drain(rc);
        CS_PIPE(cudaEventRecord(p.compute_done[k], p.compute));
        // download
        CS_PIPE(cudaStreamWaitEvent(p.out, p.compute_done[k], 0));
        if (diag_host) CS_PIPE(cudaMemcpyAsync(p.diag_pinned + 2 * i, p.diag + 2 * k, 2 * sizeof(double), cudaMemcpyDeviceToHost, p.out));
        if (copy_back)
            for (int c = 0; c < 3; ++c) CS_PIPE(cudaMemcpyAsync(hostp[c], d[c], bytes, cudaMemcpyDeviceToHost, p.out));
        CS_PIPE(cudaEventRecord(p.out_done[k], p.out));
    }
#undef CS_PIPE
    cudaError_t e = cudaStreamSynchronize(p.out);
    if (e == cudaSuccess) e = cudaStreamSynchronize(p.compute);
    if (e == cudaSuccess) e = cudaStreamSynchronize(p.in);
    if (e != cudaSuccess) return cs_fail(ctx, CS_E_CUDA, "cs_step_host_batch: %s", cudaGetErrorString(e));
    if (diag_host && members > 0) std::memcpy(diag_host, p.diag_pinned, (size_t)members * 2 * sizeof(double));
    return CS_OK;
}
