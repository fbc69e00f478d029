// K1 -- ghost-cell fill (table driven), pack / unpack for sharded contexts.
//
// One thread per ghost cell: a gather from the neighbour's interior cell, the
// seam's signed permutation of the two vector components, a scatter into the
// ghost cell.  Replaces the 12 sequential exchange_edge_pair swaps of the
// reference (JAX-DevLab-Examples.py:166-197, 238-242) with one launch; because
// every ghost cell has exactly one writer and sources are interior cells only,
// no staging and no 4-stage schedule are needed.
#include "cs_internal.cuh"

namespace {

template <typename R>
__device__ __forceinline__ void permute(R& a, R& b, int code) {
    R x = (code & 1) ? b : a;
    R y = (code & 1) ? a : b;
    a = (code & 2) ? -x : x;
    b = (code & 4) ? -y : y;
}

__device__ __forceinline__ double mul_rn(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ float mul_rn(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ double add_rn(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ float add_rn(float a, float b) { return __fadd_rn(a, b); }

// mask: bit0 scalar field h, bit1 vector field (u, v)
template <typename R>
__global__ void __launch_bounds__(256)
halo_fill_kernel(R* __restrict__ h, R* __restrict__ u, R* __restrict__ v, const int32_t* __restrict__ dst,
                 const int32_t* __restrict__ src, const uint8_t* __restrict__ code, const R* __restrict__ rot,
                 int64_t count, int mask) {
    // launched as a programmatic dependent (cs_launch_dependent): wait for the producer of h, u, v, then let the
    // consumer's blocks be scheduled while this grid runs
    cudaGridDependencySynchronize();
    cudaTriggerProgrammaticLaunchCompletion();
    for (int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k < count; k += (int64_t)gridDim.x * blockDim.x) {
        const int32_t d = dst[k], s = src[k];
        if (mask & 1) h[d] = h[s];
        if (mask & 2) {
            R a = u[s], b = v[s];
            if (rot) {   // position-dependent 2x2: two products and one sum per component, never contracted
                const R* r = rot + 4 * k;
                R x = add_rn(mul_rn(r[0], a), mul_rn(r[1], b));
                R y = add_rn(mul_rn(r[2], a), mul_rn(r[3], b));
                a = x, b = y;
            } else {
                permute(a, b, code[k]);
            }
            u[d] = a, v[d] = b;
        }
    }
}

// Interpolated ghosts (cs_halo_set_interp): ghost = ((w0 s0 + w1 s1) + w2 s2) + w3 s3 over four interior cells of the
// neighbour's row at the same depth, vectors then through a 2 x 2; every product and sum is rounded on its own (no FMA
// contraction), so a CPU evaluation in the same order gives the same bits.
template <typename R>
__device__ __forceinline__ R interp4(const R* __restrict__ f, const int4 s, const R w0, const R w1, const R w2, const R w3) {
    return add_rn(add_rn(add_rn(mul_rn(w0, f[s.x]), mul_rn(w1, f[s.y])), mul_rn(w2, f[s.z])), mul_rn(w3, f[s.w]));
}

template <typename R>
__global__ void __launch_bounds__(256)
halo_interp_kernel(R* h, R* u, R* v, const int32_t* __restrict__ dst, const int4* __restrict__ src, const R* __restrict__ w,
                   const R* __restrict__ rot, int64_t count, int mask) {
    cudaGridDependencySynchronize();
    cudaTriggerProgrammaticLaunchCompletion();
    for (int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k < count; k += (int64_t)gridDim.x * blockDim.x) {
        const int32_t d = dst[k];
        const int4 s = src[k];
        const R w0 = w[4 * k], w1 = w[4 * k + 1], w2 = w[4 * k + 2], w3 = w[4 * k + 3];
        if (mask & 1) h[d] = interp4<R>(h, s, w0, w1, w2, w3);
        if (mask & 2) {
            const R a = interp4<R>(u, s, w0, w1, w2, w3), b = interp4<R>(v, s, w0, w1, w2, w3);
            const R* r = rot + 4 * k;
            u[d] = add_rn(mul_rn(r[0], a), mul_rn(r[1], b));
            v[d] = add_rn(mul_rn(r[2], a), mul_rn(r[3], b));
        }
    }
}

// sendbuf layout: for each peer block [field][cell]; the caller slices it per peer
template <typename R>
__global__ void __launch_bounds__(256)
halo_pack_kernel(const R* __restrict__ h, const R* __restrict__ u, const R* __restrict__ v,
                 const int32_t* __restrict__ src, R* __restrict__ buf, const int64_t* __restrict__ peer_off,
                 int n_peers, int64_t count, int nfields) {
    for (int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k < count; k += (int64_t)gridDim.x * blockDim.x) {
        int r = 0;
        while (r + 1 < n_peers && k >= peer_off[r + 1]) ++r;
        const int64_t cnt = peer_off[r + 1] - peer_off[r];
        const int64_t base = peer_off[r] * nfields + (k - peer_off[r]);
        const int32_t s = src[k];
        buf[base] = h[s];
        if (nfields == 3) {
            buf[base + cnt] = u[s];
            buf[base + 2 * cnt] = v[s];
        }
    }
}

template <typename R>
__global__ void __launch_bounds__(256)
halo_unpack_kernel(R* __restrict__ h, R* __restrict__ u, R* __restrict__ v, const int32_t* __restrict__ dst,
                   const uint8_t* __restrict__ code, const R* __restrict__ buf, const int64_t* __restrict__ peer_off,
                   int n_peers, int64_t count, int nfields) {
    for (int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k < count; k += (int64_t)gridDim.x * blockDim.x) {
        int r = 0;
        while (r + 1 < n_peers && k >= peer_off[r + 1]) ++r;
        const int64_t cnt = peer_off[r + 1] - peer_off[r];
        const int64_t base = peer_off[r] * nfields + (k - peer_off[r]);
        const int32_t d = dst[k];
        h[d] = buf[base];
        if (nfields == 3) {
            R a = buf[base + cnt], b = buf[base + 2 * cnt];
            permute(a, b, code[k]);
            u[d] = a, v[d] = b;
        }
    }
}

int grid_for(const cs_ctx* ctx, int64_t count) {
    int64_t blocks = (count + 255) / 256;
    int64_t cap = (int64_t)ctx->num_sms * 8;
    return (int)(blocks < 1 ? 1 : blocks > cap ? cap : blocks);
}

}  // namespace

int cs_launch_halo(cs_ctx* ctx, void* h, void* u, void* v, int mask, cudaStream_t s) {
    if (ctx->interp_sharded) return CS_OK;     // every ghost cell of the shard is produced by cs_halo_pull (local sources too)
    if (ctx->d_interp_dst) {
        const dim3 grid(grid_for(ctx, ctx->n_interp)), block(256);
        if (ctx->dtype == CS_F64)
            CS_CUDA(ctx, cs_launch_dependent(halo_interp_kernel<double>, grid, block, 0, s, ctx->pdl, (double*)h, (double*)u, (double*)v,
                                             ctx->d_interp_dst, ctx->d_interp_src, (const double*)ctx->d_interp_w,
                                             (const double*)ctx->d_interp_rot, (int64_t)ctx->n_interp, mask));
        else
            CS_CUDA(ctx, cs_launch_dependent(halo_interp_kernel<float>, grid, block, 0, s, ctx->pdl, (float*)h, (float*)u, (float*)v,
                                             ctx->d_interp_dst, ctx->d_interp_src, (const float*)ctx->d_interp_w,
                                             (const float*)ctx->d_interp_rot, (int64_t)ctx->n_interp, mask));
        ctx->launches++;
        return CS_OK;
    }
    if (ctx->n_local == 0) return CS_OK;
    const dim3 grid(grid_for(ctx, ctx->n_local)), block(256);
    if (ctx->dtype == CS_F64)
        CS_CUDA(ctx, cs_launch_dependent(halo_fill_kernel<double>, grid, block, 0, s, ctx->pdl, (double*)h, (double*)u, (double*)v,
                                         ctx->d_halo_dst, ctx->d_halo_src, ctx->d_halo_code, (const double*)ctx->d_rot,
                                         (int64_t)ctx->n_local, mask));
    else
        CS_CUDA(ctx, cs_launch_dependent(halo_fill_kernel<float>, grid, block, 0, s, ctx->pdl, (float*)h, (float*)u, (float*)v,
                                         ctx->d_halo_dst, ctx->d_halo_src, ctx->d_halo_code, (const float*)ctx->d_rot,
                                         (int64_t)ctx->n_local, mask));
    ctx->launches++;
    return CS_OK;
}

extern "C" int cs_halo_fill_scalar(cs_ctx* ctx, void* field, void* stream) {
    if (!ctx || !field) return cs_fail(ctx, CS_E_INVALID, "cs_halo_fill_scalar: NULL argument");
    CS_ON_DEVICE(ctx, ctx->device);
    return cs_launch_halo(ctx, field, nullptr, nullptr, 1, (cudaStream_t)stream);
}

extern "C" int cs_halo_fill_vector(cs_ctx* ctx, void* u, void* v, void* stream) {
    if (!ctx || !u || !v) return cs_fail(ctx, CS_E_INVALID, "cs_halo_fill_vector: NULL argument");
    if (u == v) return cs_fail(ctx, CS_E_INVALID, "cs_halo_fill_vector: u and v must be distinct arrays");
    CS_ON_DEVICE(ctx, ctx->device);
    return cs_launch_halo(ctx, nullptr, u, v, 2, (cudaStream_t)stream);
}

extern "C" int cs_halo_fill_state(cs_ctx* ctx, void* h, void* u, void* v, void* stream) {
    if (!ctx || !h || !u || !v) return cs_fail(ctx, CS_E_INVALID, "cs_halo_fill_state: NULL argument");
    CS_ON_DEVICE(ctx, ctx->device);
    return cs_launch_halo(ctx, h, u, v, 3, (cudaStream_t)stream);
}

extern "C" int cs_halo_table_info(const cs_ctx* ctx, size_t cap, int64_t* dst_offset, int64_t* src_offset, int32_t* code,
                                  size_t* n_ghost) {
    if (!ctx) return cs_fail(ctx, CS_E_INVALID, "cs_halo_table_info: NULL context");
    if (n_ghost) *n_ghost = ctx->halo.size();
    // local entries first, in the order the fill kernel (and a rotation table) uses, then the remote ones
    size_t k = 0;
    for (int pass = 0; pass < 2; ++pass)
        for (const HaloEntry& e : ctx->halo) {
            if ((e.src >= 0) != (pass == 0)) continue;
            if (k < cap) {
                if (dst_offset) dst_offset[k] = e.dst;
                if (src_offset) src_offset[k] = e.src;
                if (code) code[k] = e.code;
            }
            ++k;
        }
    return CS_OK;
}

extern "C" int cs_halo_set_rotation(cs_ctx* ctx, const double* rot, size_t n_ghost) {
    if (!ctx) return cs_fail(ctx, CS_E_INVALID, "cs_halo_set_rotation: NULL context");
    CS_ON_DEVICE(ctx, ctx->device);
    CS_CUDA(ctx, cudaDeviceSynchronize());
    cs_drop_graphs(ctx);
    if (ctx->d_rot) cudaFree(ctx->d_rot), ctx->d_rot = nullptr;
    if (!rot) return CS_OK;
    if (n_ghost != (size_t)ctx->n_local)
        return cs_fail(ctx, CS_E_INVALID, "rotation table has %zu entries, the context has %lld local ghost cells", n_ghost,
                       (long long)ctx->n_local);
    const size_t count = n_ghost * 4;
    if (cudaMalloc(&ctx->d_rot, count * ctx->esize) != cudaSuccess) return cs_fail(ctx, CS_E_NOMEM, "cudaMalloc of rotation table failed");
    if (ctx->dtype == CS_F64) {
        CS_CUDA(ctx, cudaMemcpy(ctx->d_rot, rot, count * 8, cudaMemcpyHostToDevice));
    } else {
        std::vector<float> tmp(rot, rot + count);
        CS_CUDA(ctx, cudaMemcpy(ctx->d_rot, tmp.data(), count * 4, cudaMemcpyHostToDevice));
    }
    return CS_OK;
}

extern "C" int cs_halo_set_interp(cs_ctx* ctx, size_t n_ghost, const int64_t* dst, const int64_t* src, const double* w,
                                  const double* rot, const double* tangents) {
    if (!ctx) return cs_fail(ctx, CS_E_INVALID, "cs_halo_set_interp: NULL context");
    CS_ON_DEVICE(ctx, ctx->device);
    CS_CUDA(ctx, cudaDeviceSynchronize());
    cs_drop_graphs(ctx);
    for (void** q : {(void**)&ctx->d_interp_dst, (void**)&ctx->d_interp_src, &ctx->d_interp_w, &ctx->d_interp_rot})
        if (*q) cudaFree(*q), *q = nullptr;
    ctx->n_interp = 0;
    const int N = ctx->N, g = ctx->g;
    if (!dst) {                                           // back to the plain copy and the periodic metric
        for (int I = 0; I < N + 2 * g; ++I) ctx->xt[I] = ctx->x_base[((I - g) % N + N) % N];
        return cs_upload_metric(ctx);
    }
    if (!src || !w || !rot) return cs_fail(ctx, CS_E_INVALID, "cs_halo_set_interp: NULL table");
    if (ctx->T != ctx->T_all)
        return cs_fail(ctx, CS_E_UNSUPPORTED, "cs_halo_set_interp: interpolated ghosts need a context that owns every tile "
                                              "(the four sources of a ghost cell would live on other devices)");
    if (n_ghost != ctx->halo.size())
        return cs_fail(ctx, CS_E_INVALID, "cs_halo_set_interp: %zu rows, the context has %zu ghost cells", n_ghost, ctx->halo.size());
    const int64_t elems = (int64_t)ctx->field_elems;
    std::vector<int32_t> d32(n_ghost);
    std::vector<int4> s4(n_ghost);
    for (size_t k = 0; k < n_ghost; ++k) {
        const int64_t* q = src + 4 * k;
        if (dst[k] < 0 || dst[k] >= elems || q[0] < 0 || q[0] >= elems || q[1] < 0 || q[1] >= elems || q[2] < 0 || q[2] >= elems ||
            q[3] < 0 || q[3] >= elems)
            return cs_fail(ctx, CS_E_INVALID, "cs_halo_set_interp: row %zu has an offset outside the field", k);
        d32[k] = (int32_t)dst[k];
        s4[k] = make_int4((int)q[0], (int)q[1], (int)q[2], (int)q[3]);
    }
    const size_t count = n_ghost * 4;
    if (cudaMalloc((void**)&ctx->d_interp_dst, n_ghost * sizeof(int32_t)) != cudaSuccess ||
        cudaMalloc((void**)&ctx->d_interp_src, n_ghost * sizeof(int4)) != cudaSuccess ||
        cudaMalloc(&ctx->d_interp_w, count * ctx->esize) != cudaSuccess || cudaMalloc(&ctx->d_interp_rot, count * ctx->esize) != cudaSuccess)
        return cs_fail(ctx, CS_E_NOMEM, "cs_halo_set_interp: cudaMalloc failed");
    CS_CUDA(ctx, cudaMemcpy(ctx->d_interp_dst, d32.data(), n_ghost * sizeof(int32_t), cudaMemcpyHostToDevice));
    CS_CUDA(ctx, cudaMemcpy(ctx->d_interp_src, s4.data(), n_ghost * sizeof(int4), cudaMemcpyHostToDevice));
    if (ctx->dtype == CS_F64) {
        CS_CUDA(ctx, cudaMemcpy(ctx->d_interp_w, w, count * 8, cudaMemcpyHostToDevice));
        CS_CUDA(ctx, cudaMemcpy(ctx->d_interp_rot, rot, count * 8, cudaMemcpyHostToDevice));
    } else {
        std::vector<float> tw(w, w + count), tr(rot, rot + count);
        CS_CUDA(ctx, cudaMemcpy(ctx->d_interp_w, tw.data(), count * 4, cudaMemcpyHostToDevice));
        CS_CUDA(ctx, cudaMemcpy(ctx->d_interp_rot, tr.data(), count * 4, cudaMemcpyHostToDevice));
    }
    ctx->n_interp = (int64_t)n_ghost;
    if (tangents) {
        for (int I = 0; I < N + 2 * g; ++I) ctx->xt[I] = tangents[I];
        return cs_upload_metric(ctx);
    }
    return CS_OK;
}

namespace {
int upload_offsets(cs_ctx* ctx, bool send, int64_t** d_off);
}

extern "C" int cs_halo_interp_sharded(cs_ctx* ctx, int n_peers, const int64_t* send_counts, const int64_t* send_src,
                                      const int64_t* recv_counts, size_t n_rows, const int64_t* dst, const int64_t* src, const double* w,
                                      const double* rot, const double* tangents) {
    if (!ctx || !send_counts || !recv_counts || !dst || !src || !w || !rot)
        return cs_fail(ctx, CS_E_INVALID, "cs_halo_interp_sharded: NULL argument");
    if (n_peers < 1) return cs_fail(ctx, CS_E_INVALID, "cs_halo_interp_sharded: need at least one peer");
    if (n_rows != ctx->halo.size())
        return cs_fail(ctx, CS_E_INVALID, "cs_halo_interp_sharded: %zu rows, the context has %zu ghost cells", n_rows, ctx->halo.size());
    CS_ON_DEVICE(ctx, ctx->device);
    CS_CUDA(ctx, cudaDeviceSynchronize());
    cs_drop_graphs(ctx);
    // the exchange plan: raw cells, peer blocks in rank order
    ctx->peers.assign(n_peers, PeerPlan());
    int64_t so = 0, ro = 0;
    for (int r = 0; r < n_peers; ++r) {
        if (send_counts[r] < 0 || recv_counts[r] < 0) return cs_fail(ctx, CS_E_INVALID, "cs_halo_interp_sharded: negative count");
        ctx->peers[r].send_count = send_counts[r], ctx->peers[r].recv_count = recv_counts[r];
        ctx->peers[r].send_offset = so, ctx->peers[r].recv_offset = ro;
        so += send_counts[r], ro += recv_counts[r];
    }
    ctx->send_total = so, ctx->recv_total = ro;
    const int64_t elems = (int64_t)ctx->field_elems;
    if (elems + ro >= 2147483647LL) return cs_fail(ctx, CS_E_UNSUPPORTED, "cs_halo_interp_sharded: shard + receive buffer exceed 32-bit offsets");
    if (so > 0 && !send_src) return cs_fail(ctx, CS_E_INVALID, "cs_halo_interp_sharded: NULL send list");
    std::vector<int32_t> ss((size_t)so);
    for (int64_t k = 0; k < so; ++k) {
        if (send_src[k] < 0 || send_src[k] >= elems) return cs_fail(ctx, CS_E_INVALID, "cs_halo_interp_sharded: send offset outside the shard");
        ss[(size_t)k] = (int32_t)send_src[k];
    }
    for (void** q : {(void**)&ctx->d_send_src, (void**)&ctx->d_interp_dst, (void**)&ctx->d_interp_src, &ctx->d_interp_w, &ctx->d_interp_rot})
        if (*q) cudaFree(*q), *q = nullptr;
    if (so > 0) {
        if (cudaMalloc((void**)&ctx->d_send_src, (size_t)so * sizeof(int32_t)) != cudaSuccess)
            return cs_fail(ctx, CS_E_NOMEM, "cs_halo_interp_sharded: cudaMalloc failed");
        CS_CUDA(ctx, cudaMemcpy(ctx->d_send_src, ss.data(), (size_t)so * sizeof(int32_t), cudaMemcpyHostToDevice));
    }
    int rc;
    if ((rc = upload_offsets(ctx, true, &ctx->d_send_off))) return rc;
    if ((rc = upload_offsets(ctx, false, &ctx->d_recv_off))) return rc;
    // the rows
    std::vector<int32_t> d32(n_rows);
    std::vector<int4> s4(n_rows);
    for (size_t k = 0; k < n_rows; ++k) {
        const int64_t* q = src + 4 * k;
        if (dst[k] < 0 || dst[k] >= elems) return cs_fail(ctx, CS_E_INVALID, "cs_halo_interp_sharded: row %zu writes outside the shard", k);
        for (int j = 0; j < 4; ++j)
            if (q[j] < 0 || q[j] >= elems + ro) return cs_fail(ctx, CS_E_INVALID, "cs_halo_interp_sharded: row %zu reads outside shard + receive buffer", k);
        d32[k] = (int32_t)dst[k];
        s4[k] = make_int4((int)q[0], (int)q[1], (int)q[2], (int)q[3]);
    }
    const size_t count = n_rows * 4;
    if (cudaMalloc((void**)&ctx->d_interp_dst, n_rows * sizeof(int32_t)) != cudaSuccess ||
        cudaMalloc((void**)&ctx->d_interp_src, n_rows * sizeof(int4)) != cudaSuccess ||
        cudaMalloc(&ctx->d_interp_w, count * ctx->esize) != cudaSuccess || cudaMalloc(&ctx->d_interp_rot, count * ctx->esize) != cudaSuccess)
        return cs_fail(ctx, CS_E_NOMEM, "cs_halo_interp_sharded: cudaMalloc failed");
    CS_CUDA(ctx, cudaMemcpy(ctx->d_interp_dst, d32.data(), n_rows * sizeof(int32_t), cudaMemcpyHostToDevice));
    CS_CUDA(ctx, cudaMemcpy(ctx->d_interp_src, s4.data(), n_rows * sizeof(int4), cudaMemcpyHostToDevice));
    if (ctx->dtype == CS_F64) {
        CS_CUDA(ctx, cudaMemcpy(ctx->d_interp_w, w, count * 8, cudaMemcpyHostToDevice));
        CS_CUDA(ctx, cudaMemcpy(ctx->d_interp_rot, rot, count * 8, cudaMemcpyHostToDevice));
    } else {
        std::vector<float> tw(w, w + count), tr(rot, rot + count);
        CS_CUDA(ctx, cudaMemcpy(ctx->d_interp_w, tw.data(), count * 4, cudaMemcpyHostToDevice));
        CS_CUDA(ctx, cudaMemcpy(ctx->d_interp_rot, tr.data(), count * 4, cudaMemcpyHostToDevice));
    }
    ctx->n_interp = (int64_t)n_rows;
    ctx->interp_sharded = true;
    if (tangents) {
        for (int I = 0; I < ctx->N + 2 * ctx->g; ++I) ctx->xt[I] = tangents[I];
        return cs_upload_metric(ctx);
    }
    return CS_OK;
}

// ---------------------------------------------------------------------------------------------
// sharded exchange
// ---------------------------------------------------------------------------------------------

namespace {
int upload_offsets(cs_ctx* ctx, bool send, int64_t** d_off) {
    std::vector<int64_t> off(ctx->peers.size() + 1, 0);
    for (size_t r = 0; r < ctx->peers.size(); ++r)
        off[r + 1] = off[r] + (send ? ctx->peers[r].send_count : ctx->peers[r].recv_count);
    if (*d_off) cudaFree(*d_off), *d_off = nullptr;
    if (cudaMalloc((void**)d_off, off.size() * sizeof(int64_t)) != cudaSuccess)
        return cs_fail(ctx, CS_E_NOMEM, "cudaMalloc of peer offsets failed");
    CS_CUDA(ctx, cudaMemcpy(*d_off, off.data(), off.size() * sizeof(int64_t), cudaMemcpyHostToDevice));
    return CS_OK;
}
}  // namespace

extern "C" int cs_halo_plan(cs_ctx* ctx, int n_peers, const int32_t* peer_tile_begin, const int32_t* peer_tile_count,
                            int64_t* send_counts, int64_t* recv_counts) {
    if (!ctx) return cs_fail(ctx, CS_E_INVALID, "cs_halo_plan: NULL context");
    CS_ON_DEVICE(ctx, ctx->device);
    int rc = cs_build_peer_plan(ctx, n_peers, peer_tile_begin, peer_tile_count);
    if (rc) return rc;
    if ((rc = upload_offsets(ctx, true, &ctx->d_send_off))) return rc;
    if ((rc = upload_offsets(ctx, false, &ctx->d_recv_off))) return rc;
    for (int r = 0; r < n_peers; ++r) {
        if (send_counts) send_counts[r] = ctx->peers[r].send_count;
        if (recv_counts) recv_counts[r] = ctx->peers[r].recv_count;
    }
    return CS_OK;
}

int cs_launch_pack(cs_ctx* ctx, const void* h, const void* u, const void* v, int nfields, void* sendbuf, cudaStream_t s) {
    if (ctx->send_total == 0) return CS_OK;
    const int np = (int)ctx->peers.size();
    if (ctx->dtype == CS_F64)
        halo_pack_kernel<double><<<grid_for(ctx, ctx->send_total), 256, 0, s>>>((const double*)h, (const double*)u, (const double*)v,
                                                                               ctx->d_send_src, (double*)sendbuf, ctx->d_send_off, np,
                                                                               ctx->send_total, nfields);
    else
        halo_pack_kernel<float><<<grid_for(ctx, ctx->send_total), 256, 0, s>>>((const float*)h, (const float*)u, (const float*)v,
                                                                              ctx->d_send_src, (float*)sendbuf, ctx->d_send_off, np,
                                                                              ctx->send_total, nfields);
    ctx->launches++;
    CS_CUDA(ctx, cudaGetLastError());
    return CS_OK;
}

int cs_launch_unpack(cs_ctx* ctx, const void* recvbuf, void* h, void* u, void* v, int nfields, cudaStream_t s) {
    if (ctx->recv_total == 0) return CS_OK;
    const int np = (int)ctx->peers.size();
    if (ctx->dtype == CS_F64)
        halo_unpack_kernel<double><<<grid_for(ctx, ctx->recv_total), 256, 0, s>>>((double*)h, (double*)u, (double*)v, ctx->d_recv_dst,
                                                                                 ctx->d_recv_code, (const double*)recvbuf, ctx->d_recv_off,
                                                                                 np, ctx->recv_total, nfields);
    else
        halo_unpack_kernel<float><<<grid_for(ctx, ctx->recv_total), 256, 0, s>>>((float*)h, (float*)u, (float*)v, ctx->d_recv_dst,
                                                                                ctx->d_recv_code, (const float*)recvbuf, ctx->d_recv_off, np,
                                                                                ctx->recv_total, nfields);
    ctx->launches++;
    CS_CUDA(ctx, cudaGetLastError());
    return CS_OK;
}

extern "C" int cs_halo_pack(cs_ctx* ctx, const void* h, const void* u, const void* v, int nfields, void* sendbuf, void* stream) {
    if (!ctx || !h || !sendbuf || (nfields == 3 && (!u || !v))) return cs_fail(ctx, CS_E_INVALID, "cs_halo_pack: NULL argument");
    if (nfields != 1 && nfields != 3) return cs_fail(ctx, CS_E_INVALID, "cs_halo_pack: nfields must be 1 or 3");
    if (ctx->peers.empty()) return cs_fail(ctx, CS_E_INVALID, "cs_halo_pack: call cs_halo_plan first");
    CS_ON_DEVICE(ctx, ctx->device);
    return cs_launch_pack(ctx, h, u, v, nfields, sendbuf, (cudaStream_t)stream);
}

extern "C" int cs_halo_unpack(cs_ctx* ctx, const void* recvbuf, void* h, void* u, void* v, int nfields, void* stream) {
    if (!ctx || !h || !recvbuf || (nfields == 3 && (!u || !v))) return cs_fail(ctx, CS_E_INVALID, "cs_halo_unpack: NULL argument");
    if (nfields != 1 && nfields != 3) return cs_fail(ctx, CS_E_INVALID, "cs_halo_unpack: nfields must be 1 or 3");
    if (ctx->peers.empty()) return cs_fail(ctx, CS_E_INVALID, "cs_halo_unpack: call cs_halo_plan first");
    CS_ON_DEVICE(ctx, ctx->device);
    return cs_launch_unpack(ctx, recvbuf, h, u, v, nfields, (cudaStream_t)stream);
}
