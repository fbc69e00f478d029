// One-sided halo exchange between the GPUs of one box: peer stores over NVLink instead of pack -> NCCL -> unpack.
//
// Every context owns a double-buffered receive buffer and a flag word per source rank (cudaMalloc'd here, exported
// through CUDA IPC handles and mapped by the peers once, at set-up).  A ghost fill is then two launches:
//   cs_halo_push  gathers the boundary cells each peer needs and stores them STRAIGHT INTO THAT PEER'S receive buffer
//                 (plain st.global through the peer mapping); the last block to finish makes the stores visible
//                 system-wide and writes the fill's epoch into the peer's flag word for this rank;
//   cs_halo_pull  waits until every source rank's flag has reached the epoch, then scatters the receive buffer into
//                 the ghost cells with the seam's signed permutation (as cs_halo_unpack does).
// Epochs count fills (one push and one pull per fill on every rank, kept in device memory so that a replayed CUDA
// graph needs no new arguments); buffer parity = epoch & 1.  Two buffers suffice: a rank's push for fill e + 2 comes
// after its own pull of fill e + 1, which needs the peer's push e + 1, which the peer's stream ordered after its
// pull of fill e -- so nobody overwrites a buffer that is still being read.
// Replaces: the cross-device copies GSPMD inserts for .at[face].set(...) (JAX-DevLab-Examples.py:79,192-195); the
// north_star lists CUDA-IPC peer stores first, NCCL second (cs_halo_pack / unpack + all-to-all stays as that path).
#include <cstring>

#include "cs_internal.cuh"

struct cs_peer_state {
    void* recv[2] = {nullptr, nullptr};          // own receive buffers, recv_total * 3 elements each
    unsigned int* flags = nullptr;               // own flags [2][n_peers] (parity, source rank), then 4 counters
    std::vector<void*> opened;                   // IPC mappings to close
    void** d_peer_recv = nullptr;                // [2][n_peers] peers' receive buffers as seen from here
    unsigned int** d_peer_flags = nullptr;       // [n_peers] peers' flag arrays as seen from here
    int64_t* d_remote_off = nullptr;             // [n_peers] where this rank's cells start in peer r's buffer (cells)
    int me = -1, n_peers = 0;
    bool connected = false;
};

namespace {

constexpr int PEER_THREADS = 256;

__device__ __forceinline__ void st_release_sys(unsigned int* p, unsigned int v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int ld_acquire_sys(const unsigned int* p) {
    unsigned int v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// counters (device, own): [0] pushes done, [1] pulls done, [2] push ticket, [3] pull ticket, [4] waits that timed out
template <typename R>
__global__ void __launch_bounds__(PEER_THREADS)
halo_push_kernel(const R* __restrict__ h, const R* __restrict__ u, const R* __restrict__ v, const int32_t* __restrict__ src,
                 void* const* __restrict__ peer_recv, unsigned int* const* __restrict__ peer_flags,
                 const int64_t* __restrict__ peer_off, const int64_t* __restrict__ remote_off, int n_peers, int me, int64_t count,
                 int nfields, int aos, unsigned int* counters) {
    const unsigned int epoch = counters[0] + 1u;          // written only by the last block of the previous push
    const int parity = (int)(epoch & 1u);
    for (int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k < count; k += (int64_t)gridDim.x * blockDim.x) {
        int r = 0;
        while (r + 1 < n_peers && k >= peer_off[r + 1]) ++r;
        const int64_t cnt = peer_off[r + 1] - peer_off[r];
        R* buf = static_cast<R*>(peer_recv[parity * n_peers + r]);       // the PEER's memory
        const int32_t s = src[k];
        if (aos) {                                        // [cell][field]: the receiver interpolates raw cells (second-order ghosts)
            const int64_t base = (remote_off[r] + (k - peer_off[r])) * nfields;
            buf[base] = h[s];
            if (nfields == 3) buf[base + 1] = u[s], buf[base + 2] = v[s];
        } else {                                          // per source rank [field][cell]
            const int64_t base = remote_off[r] * nfields + (k - peer_off[r]);
            buf[base] = h[s];
            if (nfields == 3) {
                buf[base + cnt] = u[s];
                buf[base + 2 * cnt] = v[s];
            }
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence_system();                           // this block's stores are visible to the peers ...
        if (atomicAdd(&counters[2], 1u) == gridDim.x - 1) {
            __threadfence_system();                       // ... and so are everybody else's: tell the peers
            for (int r = 0; r < n_peers; ++r)
                if (r != me && peer_off[r + 1] > peer_off[r]) st_release_sys(peer_flags[r] + parity * n_peers + me, epoch);
            counters[2] = 0;
            counters[0] = epoch;
        }
    }
}

// Wait (thread 0 of every block) until every source rank's flag for this fill has arrived.  Epochs wrap: compared as a
// signed difference.  A peer that never issues this fill (a rank that died, or a caller that did not keep the ranks'
// fills in step) must not hang the GPU: after `timeout` SM cycles the wait gives up and counts the event
// (cs_halo_peer_timeouts).  The failure is sticky -- once a wait has timed out no later one waits at all, so a broken
// job drains at full speed instead of paying the timeout on every fill -- and the host mirror raises on it.
__device__ __forceinline__ void wait_for_peers(const unsigned int* flags, const int64_t* peer_off, int n_peers, int me, int parity,
                                               unsigned int epoch, unsigned int* counters, long long timeout) {
    if (threadIdx.x == 0 && *(volatile unsigned int*)&counters[4] == 0u) {
        for (int r = 0; r < n_peers; ++r) {
            if (r == me || peer_off[r + 1] == peer_off[r]) continue;
            const unsigned int* f = flags + parity * n_peers + r;
            const long long t_start = clock64();
            bool gave_up = false;
            while ((int)(ld_acquire_sys(f) - epoch) < 0) {
                __nanosleep(64);
                if (clock64() - t_start > timeout || *(volatile unsigned int*)&counters[4] != 0u) {
                    atomicAdd(&counters[4], 1u);
                    gave_up = true;
                    break;
                }
            }
            if (gave_up) break;
        }
    }
    __syncthreads();
}

__device__ __forceinline__ double mul_rn(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ float mul_rn(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ double add_rn(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ float add_rn(float a, float b) { return __fadd_rn(a, b); }

// Second-order ghosts on a shard: every ghost cell = ((w0 s0 + w1 s1) + w2 s2) + w3 s3 (then the 2 x 2 for vectors), the
// sources addressed in one virtual space -- offsets below field_elems are cells of the local shard, field_elems + k is
// cell k of the receive buffer ([cell][field], what the peers' pushes wrote).  Same rounding as halo_interp_kernel.
template <typename R>
__global__ void __launch_bounds__(PEER_THREADS)
halo_interp_pull_kernel(R* h, R* u, R* v, const int32_t* __restrict__ dst, const int4* __restrict__ src, const R* __restrict__ w,
                        const R* __restrict__ rot, const R* recv0, const R* recv1, const unsigned int* flags,
                        const int64_t* __restrict__ peer_off, int n_peers, int me, int64_t rows, int field_elems, int nfields,
                        unsigned int* counters, long long timeout) {
    const unsigned int epoch = counters[1] + 1u;
    const int parity = (int)(epoch & 1u);
    wait_for_peers(flags, peer_off, n_peers, me, parity, epoch, counters, timeout);
    const R* recv = parity ? recv1 : recv0;
    auto get = [&](const R* f, int fi, int o) -> R {
        return o < field_elems ? f[o] : __ldcg(recv + (size_t)(o - field_elems) * nfields + fi);
    };
    auto interp = [&](const R* f, int fi, const int4 s, R w0, R w1, R w2, R w3) -> R {
        return add_rn(add_rn(add_rn(mul_rn(w0, get(f, fi, s.x)), mul_rn(w1, get(f, fi, s.y))), mul_rn(w2, get(f, fi, s.z))),
                      mul_rn(w3, get(f, fi, s.w)));
    };
    for (int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k < rows; k += (int64_t)gridDim.x * blockDim.x) {
        const int32_t d = dst[k];
        const int4 s = src[k];
        const R w0 = w[4 * k], w1 = w[4 * k + 1], w2 = w[4 * k + 2], w3 = w[4 * k + 3];
        h[d] = interp(h, 0, s, w0, w1, w2, w3);
        if (nfields == 3) {
            const R a = interp(u, 1, s, w0, w1, w2, w3), b = interp(v, 2, s, w0, w1, w2, w3);
            const R* r = rot + 4 * k;
            u[d] = add_rn(mul_rn(r[0], a), mul_rn(r[1], b));
            v[d] = add_rn(mul_rn(r[2], a), mul_rn(r[3], b));
        }
    }
    __syncthreads();
    if (threadIdx.x == 0 && atomicAdd(&counters[3], 1u) == gridDim.x - 1) {
        counters[3] = 0;
        counters[1] = epoch;
    }
}

template <typename R>
__global__ void __launch_bounds__(PEER_THREADS)
halo_pull_kernel(R* __restrict__ h, R* __restrict__ u, R* __restrict__ v, const int32_t* __restrict__ dst,
                 const uint8_t* __restrict__ code, const R* recv0, const R* recv1, const unsigned int* flags,
                 const int64_t* __restrict__ peer_off, int n_peers, int me, int64_t count, int nfields, unsigned int* counters,
                 long long timeout) {
    const unsigned int epoch = counters[1] + 1u;
    const int parity = (int)(epoch & 1u);
    wait_for_peers(flags, peer_off, n_peers, me, parity, epoch, counters, timeout);   // on a time-out the scatter goes on with stale data
    const R* buf = parity ? recv1 : recv0;
    for (int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k < count; k += (int64_t)gridDim.x * blockDim.x) {
        int r = 0;
        while (r + 1 < n_peers && k >= peer_off[r + 1]) ++r;
        const int64_t cnt = peer_off[r + 1] - peer_off[r];
        const int64_t base = peer_off[r] * nfields + (k - peer_off[r]);
        const int32_t d = dst[k];
        // the peers wrote this memory: read it through L2 (ld.cg), never from a line L1 may still hold from the
        // fill two epochs ago
        h[d] = __ldcg(buf + base);
        if (nfields == 3) {
            R a = __ldcg(buf + base + cnt), b = __ldcg(buf + base + 2 * cnt);
            const int c = code[k];
            const R x = (c & 1) ? b : a, y = (c & 1) ? a : b;
            u[d] = (c & 2) ? -x : x;
            v[d] = (c & 4) ? -y : y;
        }
    }
    __syncthreads();
    if (threadIdx.x == 0 && atomicAdd(&counters[3], 1u) == gridDim.x - 1) {
        counters[3] = 0;
        counters[1] = epoch;
    }
}

int grid_for(const cs_ctx* ctx, int64_t count) {
    int64_t blocks = (count + PEER_THREADS - 1) / PEER_THREADS;
    const int64_t cap = (int64_t)ctx->num_sms * 2;        // a pull may have to wait for a peer: keep it small
    return (int)(blocks < 1 ? 1 : blocks > cap ? cap : blocks);
}

void peer_free(cs_ctx* ctx) {
    cs_peer_state* P = ctx->peer;
    if (!P) return;
    for (void* q : P->opened) cudaIpcCloseMemHandle(q);
    for (void* q : {P->recv[0], P->recv[1], (void*)P->flags, (void*)P->d_peer_recv, (void*)P->d_peer_flags, (void*)P->d_remote_off})
        if (q) cudaFree(q);
    delete P;
    ctx->peer = nullptr;
}

}  // namespace

void cs_peer_destroy(cs_ctx* ctx) { peer_free(ctx); }

static_assert(CS_PEER_HANDLE_BYTES == 3 * sizeof(cudaIpcMemHandle_t), "handle block size");

extern "C" int cs_halo_peer_alloc(cs_ctx* ctx, void* handles_out, size_t handles_bytes) {
    if (!ctx || !handles_out) return cs_fail(ctx, CS_E_INVALID, "cs_halo_peer_alloc: NULL argument");
    if (handles_bytes < CS_PEER_HANDLE_BYTES)
        return cs_fail(ctx, CS_E_INVALID, "cs_halo_peer_alloc: handle buffer of %zu bytes, need %zu", handles_bytes, CS_PEER_HANDLE_BYTES);
    if (ctx->peers.empty()) return cs_fail(ctx, CS_E_INVALID, "cs_halo_peer_alloc: call cs_halo_plan first");
    CS_ON_DEVICE(ctx, ctx->device);
    peer_free(ctx);
    cs_peer_state* P = new cs_peer_state();
    ctx->peer = P;
    P->n_peers = (int)ctx->peers.size();
    const size_t bytes = (size_t)(ctx->recv_total > 0 ? ctx->recv_total : 1) * 3 * ctx->esize;
    const size_t fbytes = ((size_t)2 * P->n_peers + 8) * sizeof(unsigned int);
    if (cudaMalloc(&P->recv[0], bytes) != cudaSuccess || cudaMalloc(&P->recv[1], bytes) != cudaSuccess ||
        cudaMalloc((void**)&P->flags, fbytes) != cudaSuccess)
        return cs_fail(ctx, CS_E_NOMEM, "cs_halo_peer_alloc: cudaMalloc failed");
    CS_CUDA(ctx, cudaMemset(P->flags, 0, fbytes));
    cudaIpcMemHandle_t hd[3];
    CS_CUDA(ctx, cudaIpcGetMemHandle(&hd[0], P->recv[0]));
    CS_CUDA(ctx, cudaIpcGetMemHandle(&hd[1], P->recv[1]));
    CS_CUDA(ctx, cudaIpcGetMemHandle(&hd[2], P->flags));
    std::memcpy(handles_out, hd, sizeof hd);
    return CS_OK;
}

namespace {
// upload the peers' buffer addresses (as this device sees them) and the offsets of this rank's blocks in their buffers
int peer_finish(cs_ctx* ctx, cs_peer_state* P, int n_peers, int my_rank, const std::vector<void*>& recv, const std::vector<unsigned int*>& flags,
                const int64_t* remote_offsets) {
    P->me = my_rank;
    std::vector<int64_t> off(remote_offsets, remote_offsets + n_peers);
    for (void* q : {(void*)P->d_peer_recv, (void*)P->d_peer_flags, (void*)P->d_remote_off})
        if (q) cudaFree(q);
    P->d_peer_recv = nullptr, P->d_peer_flags = nullptr, P->d_remote_off = nullptr;
    if (cudaMalloc((void**)&P->d_peer_recv, recv.size() * sizeof(void*)) != cudaSuccess ||
        cudaMalloc((void**)&P->d_peer_flags, flags.size() * sizeof(unsigned int*)) != cudaSuccess ||
        cudaMalloc((void**)&P->d_remote_off, off.size() * sizeof(int64_t)) != cudaSuccess)
        return cs_fail(ctx, CS_E_NOMEM, "cs_halo_peer_connect: cudaMalloc failed");
    CS_CUDA(ctx, cudaMemcpy(P->d_peer_recv, recv.data(), recv.size() * sizeof(void*), cudaMemcpyHostToDevice));
    CS_CUDA(ctx, cudaMemcpy(P->d_peer_flags, flags.data(), flags.size() * sizeof(unsigned int*), cudaMemcpyHostToDevice));
    CS_CUDA(ctx, cudaMemcpy(P->d_remote_off, off.data(), off.size() * sizeof(int64_t), cudaMemcpyHostToDevice));
    P->connected = true;
    return CS_OK;
}
}  // namespace

extern "C" int cs_halo_peer_connect(cs_ctx* ctx, int n_peers, int my_rank, const void* all_handles, const int64_t* remote_offsets) {
    if (!ctx || !all_handles || !remote_offsets) return cs_fail(ctx, CS_E_INVALID, "cs_halo_peer_connect: NULL argument");
    cs_peer_state* P = ctx->peer;
    if (!P) return cs_fail(ctx, CS_E_INVALID, "cs_halo_peer_connect: call cs_halo_peer_alloc first");
    if (n_peers != P->n_peers || my_rank < 0 || my_rank >= n_peers)
        return cs_fail(ctx, CS_E_INVALID, "cs_halo_peer_connect: %d peers / rank %d do not match the plan (%d peers)", n_peers, my_rank, P->n_peers);
    CS_ON_DEVICE(ctx, ctx->device);
    std::vector<void*> recv(2 * (size_t)n_peers, nullptr);
    std::vector<unsigned int*> flags(n_peers, nullptr);
    const cudaIpcMemHandle_t* hd = static_cast<const cudaIpcMemHandle_t*>(all_handles);
    for (int r = 0; r < n_peers; ++r) {
        if (r == my_rank) {
            recv[r] = P->recv[0], recv[n_peers + r] = P->recv[1], flags[r] = P->flags;
            continue;
        }
        if (ctx->peers[r].send_count == 0 && ctx->peers[r].recv_count == 0) continue;      // not a neighbour
        void* q[3];
        for (int k = 0; k < 3; ++k) {
            cudaError_t e = cudaIpcOpenMemHandle(&q[k], hd[3 * r + k], cudaIpcMemLazyEnablePeerAccess);
            if (e != cudaSuccess)
                return cs_fail(ctx, CS_E_CUDA, "cs_halo_peer_connect: cudaIpcOpenMemHandle for rank %d failed: %s (peers must be GPUs of one "
                                               "box with peer access, one process per GPU)", r, cudaGetErrorString(e));
            P->opened.push_back(q[k]);
        }
        recv[r] = q[0], recv[n_peers + r] = q[1], flags[r] = static_cast<unsigned int*>(q[2]);
    }
    return peer_finish(ctx, P, n_peers, my_rank, recv, flags, remote_offsets);
}

extern "C" int cs_halo_peer_pointers(cs_ctx* ctx, void** out3) {
    if (!ctx || !out3) return cs_fail(ctx, CS_E_INVALID, "cs_halo_peer_pointers: NULL argument");
    cs_peer_state* P = ctx->peer;
    if (!P) return cs_fail(ctx, CS_E_INVALID, "cs_halo_peer_pointers: call cs_halo_peer_alloc first");
    out3[0] = P->recv[0], out3[1] = P->recv[1], out3[2] = P->flags;
    return CS_OK;
}

extern "C" int cs_halo_peer_connect_pointers(cs_ctx* ctx, int n_peers, int my_rank, void* const* all_pointers, const int64_t* remote_offsets) {
    if (!ctx || !all_pointers || !remote_offsets) return cs_fail(ctx, CS_E_INVALID, "cs_halo_peer_connect_pointers: NULL argument");
    cs_peer_state* P = ctx->peer;
    if (!P) return cs_fail(ctx, CS_E_INVALID, "cs_halo_peer_connect_pointers: call cs_halo_peer_alloc first");
    if (n_peers != P->n_peers || my_rank < 0 || my_rank >= n_peers)
        return cs_fail(ctx, CS_E_INVALID, "cs_halo_peer_connect_pointers: %d peers / rank %d do not match the plan (%d peers)", n_peers, my_rank,
                       P->n_peers);
    CS_ON_DEVICE(ctx, ctx->device);
    std::vector<void*> recv(2 * (size_t)n_peers, nullptr);
    std::vector<unsigned int*> flags(n_peers, nullptr);
    for (int r = 0; r < n_peers; ++r) {
        recv[r] = all_pointers[3 * r], recv[n_peers + r] = all_pointers[3 * r + 1];
        flags[r] = static_cast<unsigned int*>(all_pointers[3 * r + 2]);
    }
    return peer_finish(ctx, P, n_peers, my_rank, recv, flags, remote_offsets);
}

extern "C" int cs_halo_push(cs_ctx* ctx, const void* h, const void* u, const void* v, int nfields, void* stream) {
    if (!ctx || !h || (nfields == 3 && (!u || !v))) return cs_fail(ctx, CS_E_INVALID, "cs_halo_push: NULL argument");
    if (nfields != 1 && nfields != 3) return cs_fail(ctx, CS_E_INVALID, "cs_halo_push: nfields must be 1 or 3");
    cs_peer_state* P = ctx->peer;
    if (!P || !P->connected) return cs_fail(ctx, CS_E_INVALID, "cs_halo_push: call cs_halo_peer_alloc / cs_halo_peer_connect first");
    CS_ON_DEVICE(ctx, ctx->device);
    cudaStream_t s = (cudaStream_t)stream;
    unsigned int* counters = P->flags + 2 * P->n_peers;
    const int grid = grid_for(ctx, ctx->send_total);
    if (ctx->dtype == CS_F64)
        halo_push_kernel<double><<<grid, PEER_THREADS, 0, s>>>((const double*)h, (const double*)u, (const double*)v, ctx->d_send_src,
                                                              P->d_peer_recv, P->d_peer_flags, ctx->d_send_off, P->d_remote_off, P->n_peers,
                                                              P->me, ctx->send_total, nfields, ctx->interp_sharded ? 1 : 0, counters);
    else
        halo_push_kernel<float><<<grid, PEER_THREADS, 0, s>>>((const float*)h, (const float*)u, (const float*)v, ctx->d_send_src,
                                                             P->d_peer_recv, P->d_peer_flags, ctx->d_send_off, P->d_remote_off, P->n_peers,
                                                             P->me, ctx->send_total, nfields, ctx->interp_sharded ? 1 : 0, counters);
    ctx->launches++;
    CS_CUDA(ctx, cudaGetLastError());
    return CS_OK;
}

extern "C" int cs_halo_pull(cs_ctx* ctx, void* h, void* u, void* v, int nfields, void* stream) {
    if (!ctx || !h || (nfields == 3 && (!u || !v))) return cs_fail(ctx, CS_E_INVALID, "cs_halo_pull: NULL argument");
    if (nfields != 1 && nfields != 3) return cs_fail(ctx, CS_E_INVALID, "cs_halo_pull: nfields must be 1 or 3");
    cs_peer_state* P = ctx->peer;
    if (!P || !P->connected) return cs_fail(ctx, CS_E_INVALID, "cs_halo_pull: call cs_halo_peer_alloc / cs_halo_peer_connect first");
    CS_ON_DEVICE(ctx, ctx->device);
    cudaStream_t s = (cudaStream_t)stream;
    unsigned int* counters = P->flags + 2 * P->n_peers;
    // option "peer_timeout_ms" (CS_PEER_TIMEOUT_MS; default 5000) in SM cycles at the device's nominal clock
    const long long timeout = (long long)(ctx->tune.peer_timeout_ms > 0 ? ctx->tune.peer_timeout_ms : 1) * ctx->clock_khz;
    if (ctx->interp_sharded) {
        const int grid = grid_for(ctx, ctx->n_interp);
        if (ctx->dtype == CS_F64)
            halo_interp_pull_kernel<double><<<grid, PEER_THREADS, 0, s>>>((double*)h, (double*)u, (double*)v, ctx->d_interp_dst, ctx->d_interp_src,
                                                                         (const double*)ctx->d_interp_w, (const double*)ctx->d_interp_rot,
                                                                         (const double*)P->recv[0], (const double*)P->recv[1], P->flags,
                                                                         ctx->d_recv_off, P->n_peers, P->me, ctx->n_interp,
                                                                         (int)ctx->field_elems, nfields, counters, timeout);
        else
            halo_interp_pull_kernel<float><<<grid, PEER_THREADS, 0, s>>>((float*)h, (float*)u, (float*)v, ctx->d_interp_dst, ctx->d_interp_src,
                                                                        (const float*)ctx->d_interp_w, (const float*)ctx->d_interp_rot,
                                                                        (const float*)P->recv[0], (const float*)P->recv[1], P->flags,
                                                                        ctx->d_recv_off, P->n_peers, P->me, ctx->n_interp,
                                                                        (int)ctx->field_elems, nfields, counters, timeout);
        ctx->launches++;
        CS_CUDA(ctx, cudaGetLastError());
        return CS_OK;
    }
    const int grid = grid_for(ctx, ctx->recv_total);
    if (ctx->dtype == CS_F64)
        halo_pull_kernel<double><<<grid, PEER_THREADS, 0, s>>>((double*)h, (double*)u, (double*)v, ctx->d_recv_dst, ctx->d_recv_code,
                                                              (const double*)P->recv[0], (const double*)P->recv[1], P->flags, ctx->d_recv_off,
                                                              P->n_peers, P->me, ctx->recv_total, nfields, counters, timeout);
    else
        halo_pull_kernel<float><<<grid, PEER_THREADS, 0, s>>>((float*)h, (float*)u, (float*)v, ctx->d_recv_dst, ctx->d_recv_code,
                                                             (const float*)P->recv[0], (const float*)P->recv[1], P->flags, ctx->d_recv_off,
                                                             P->n_peers, P->me, ctx->recv_total, nfields, counters, timeout);
    ctx->launches++;
    CS_CUDA(ctx, cudaGetLastError());
    return CS_OK;
}

extern "C" int cs_halo_peer_timeouts(cs_ctx* ctx, unsigned int* count_host) {
    if (!ctx || !count_host) return cs_fail(ctx, CS_E_INVALID, "cs_halo_peer_timeouts: NULL argument");
    cs_peer_state* P = ctx->peer;
    if (!P) return cs_fail(ctx, CS_E_INVALID, "cs_halo_peer_timeouts: no peer exchange on this context");
    CS_ON_DEVICE(ctx, ctx->device);
    CS_CUDA(ctx, cudaMemcpy(count_host, P->flags + 2 * P->n_peers + 4, sizeof(unsigned int), cudaMemcpyDeviceToHost));
    return CS_OK;
}
