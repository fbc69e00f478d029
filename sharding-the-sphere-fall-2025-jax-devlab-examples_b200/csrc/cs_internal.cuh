// Internal declarations shared by the translation units of libcubesphere_b200.so.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <string>
#include <vector>

#include "cubesphere.h"

#define CS_MAX_HALO 4

struct HaloEntry {           // one ghost cell
    int64_t dst;             // element offset in the local shard
    int64_t src;             // element offset in the local shard, or -1 when remote
    int32_t src_tile;        // global tile id of the source
    int32_t src_local;       // element offset inside the source tile
    int32_t code;            // bit0 swap, bit1 negate out-u, bit2 negate out-v
};

struct PeerPlan {
    int32_t tile_begin = 0, tile_count = 0;
    int64_t send_count = 0, recv_count = 0;      // cells
    int64_t send_offset = 0, recv_offset = 0;    // cells, prefix sums over peers
};

struct StepGraph {            // one captured step, keyed by what is baked into its kernel parameters
    cudaGraphExec_t exec = nullptr;
    void *h = nullptr, *u = nullptr, *v = nullptr;
    double dt = 0.0;
    double* diag = nullptr;
    bool topo = false;
    int64_t launches = 0;
    uint64_t stamp = 0;
};

// Tuning knobs, read from the CS_* environment variables ONCE, when the context is created.
struct CsTuning {
    int march_prefetch = 2;      // CS_MARCH_PREFETCH: L2 prefetch distance in rows
    int march_rows = 0;          // CS_MARCH_ROWS: strip height, 0 = wave-aware heuristic
    int march_vec = 4;           // CS_MARCH_VEC: fp32 columns per thread (4 | 2)
    int march_unit = 1;          // CS_MARCH_UNIT: first-stage build (0 never, 1 fp64 + fp32 launches without sums, 2 always)
    int reserve_sms = 12;        // CS_MARCH_RESERVE_SMS: SMs left to the exchange stream by a split stage
    int tma_ring = 6;            // CS_TMA_RING
    int fused_halo = 1;          // CS_FUSED_HALO: 0 vetoes cfg.fused_halo
    int peer_timeout_ms = 5000;  // CS_PEER_TIMEOUT_MS: how long a one-sided pull waits for a peer's flag before it gives up
    int step_graph = 1;          // CS_STEP_GRAPH: 0 = cs_step enqueues its launches instead of replaying a graph
    int step_debug = 0;          // CS_STEP_DEBUG: timing experiments on the step kernel (wrong results)
    int step_kernel = 0;         // CS_STEP_KERNEL: 1 = cs_step uses the one-launch-per-step kernel where it can (measured
                                 // slower than the replayed graph of stage launches at every size: DESIGN.md 4.2)
    int fused_diag = 1;          // CS_FUSED_DIAG: 0 = diagnostics of a step by the separate reduction kernel
};

struct cs_peer_state;   // one-sided exchange over CUDA IPC mappings (cs_peer.cu)
void cs_peer_destroy(struct cs_ctx* ctx);

struct cs_ctx {
    cs_config cfg;
    cs_peer_state* peer = nullptr;
    CsTuning tune;
    int N = 0, p = 0, n = 0, g = 0, m = 0;
    int T_all = 0, t0 = 0, T = 0;                // t0, T: first POSITION along axis 0 and number of positions owned
    std::vector<int32_t> order, inv;             // position -> tile id (cfg.tile_order or the identity) and its inverse
    int32_t* d_tile_ids = nullptr;               // [T] tile ids of the owned positions
    bool plain_order = true;                     // ... which are simply t0, t0 + 1, ...: kernels skip the table
    int local_index(int tile) const {            // index of a tile inside this context's shard, or -1
        const int pos = inv[tile];
        return pos >= t0 && pos < t0 + T ? pos - t0 : -1;
    }
    int dtype = CS_F64;
    size_t esize = 8;
    int device = 0, num_sms = 148;
    size_t field_elems = 0;                      // T * m * m
    bool pitch16 = false;                        // row pitch is a multiple of 16 bytes
    double radius, omega, gravity, axis[3];
    double delta;
    double cf[6][3];                             // Coriolis: sin(lat_rot) = (c0 + c1 X + c2 Y) / rho
    std::vector<double> xt;                      // face-level periodic tangent table, N + 2g entries

    // device tables (in the state's dtype)
    void* d_xt = nullptr;                        // X at padded face-level index
    void* d_ainv = nullptr;                      // 1 / (1 + X^2)
    double* d_xt64 = nullptr;                    // fp64 copy of d_xt for the diagnostics

    // ghost cells
    std::vector<HaloEntry> halo;                 // all ghost cells of the owned tiles
    int64_t n_local = 0;                         // entries with src >= 0
    int32_t* d_halo_dst = nullptr;               // [n_local]
    int32_t* d_halo_src = nullptr;
    uint8_t* d_halo_code = nullptr;
    void* d_rot = nullptr;                       // optional [n_local][4] rotation (state dtype)
    // optional interpolated ghosts (cs_halo_set_interp): every ghost = sum of 4 weighted interior cells, then a 2 x 2
    int64_t n_interp = 0;
    int32_t* d_interp_dst = nullptr;             // [n_interp]
    int4* d_interp_src = nullptr;                // [n_interp] four source offsets
    void* d_interp_w = nullptr;                  // [n_interp][4] weights (state dtype)
    void* d_interp_rot = nullptr;                // [n_interp][4] row-major 2 x 2 (state dtype)
    bool interp_sharded = false;                 // ... on a shard: sources may be cells of the receive buffer (cs_halo_interp_sharded)
    std::vector<double> x_base;                  // the N centre tangents the context was created with
    int4* d_push = nullptr;                      // [T][4] (dst local tile or -1, dst edge, reverse, code)
    bool push_complete = false;                  // every neighbour tile is local
    int clock_khz = 1500000;                     // SM clock of the device (cudaDevAttrClockRate), for time-outs in cycles
    bool stage_pushed = false;                   // the last stage launch stored the edge ghosts itself
    bool stage_diag_fused = false;               // the last stage launch wrote its diag_out itself (K3 fused into K2)
    bool pdl = false;                            // programmatic dependent launch for stage / halo kernels (CS_PDL=1: on)

    // sharded exchange
    std::vector<PeerPlan> peers;
    int64_t send_total = 0, recv_total = 0;
    int32_t* d_send_src = nullptr;               // [send_total] element offsets to gather
    int32_t* d_recv_dst = nullptr;               // [recv_total]
    uint8_t* d_recv_code = nullptr;
    int64_t* d_send_off = nullptr;               // [n_peers + 1] prefix sums, cells
    int64_t* d_recv_off = nullptr;

    // scratch state for the RK stages and topography
    void* qa[3] = {nullptr, nullptr, nullptr};
    void* qb[3] = {nullptr, nullptr, nullptr};
    void* topo = nullptr;
    void* host_state[3] = {nullptr, nullptr, nullptr};   // device copy used by cs_step_host
    struct HostPipe {                                     // cs_step_host_batch: two device slots, three streams
        bool ready = false;
        void* state[2][3] = {{nullptr, nullptr, nullptr}, {nullptr, nullptr, nullptr}};   // slot 0 aliases host_state
        double* diag = nullptr;                           // [2][2]
        double* diag_pinned = nullptr;                    // page-locked staging for the members' [mass, energy]
        int diag_cap = 0;                                 // ... in members
        cudaStream_t in = nullptr, compute = nullptr, out = nullptr;
        cudaEvent_t in_done[2] = {nullptr, nullptr}, compute_done[2] = {nullptr, nullptr}, out_done[2] = {nullptr, nullptr};
    } pipe;
    bool has_topo = false;

    // reductions
    double* d_partials = nullptr;                // [partial_cap][2]
    int partial_cap = 0;
    double* d_diag = nullptr;                    // [2]
    unsigned int* d_ticket = nullptr;

    StepGraph graphs[4];
    uint64_t graph_clock = 0;
    cudaStream_t cap_stream = nullptr;

    mutable std::string err;
    int64_t launches = 0;
    int last_stage_kernel = CS_KERNEL_AUTO;      // variant of the most recent stage launch (cs_stage_kernel_info)
    int64_t naive_fallbacks = 0;                 // CS_KERNEL_AUTO launches that had to take the one-thread-per-cell kernel
};

// Every entry point runs on the context's device and puts the calling thread's current device back on return: a
// handler thread of another framework must not find its device switched under it.
struct CsDeviceGuard {
    int prev = -1, want = -1;
    cudaError_t err = cudaSuccess;
    explicit CsDeviceGuard(int device) : want(device) {
        err = cudaGetDevice(&prev);
        if (err != cudaSuccess) prev = -1;
        if (err == cudaSuccess && prev != device) err = cudaSetDevice(device);
    }
    ~CsDeviceGuard() {
        if (prev >= 0 && prev != want) cudaSetDevice(prev);
    }
    CsDeviceGuard(const CsDeviceGuard&) = delete;
    CsDeviceGuard& operator=(const CsDeviceGuard&) = delete;
};

#define CS_ON_DEVICE(ctx, device)                                                                                  \
    CsDeviceGuard cs_guard__(device);                                                                              \
    if (cs_guard__.err != cudaSuccess)                                                                             \
        return cs_fail(ctx, CS_E_CUDA, "selecting device %d failed: %s", (int)(device), cudaGetErrorString(cs_guard__.err))

int cs_fail(const cs_ctx* ctx, int code, const char* fmt, ...);
void cs_drop_graphs(cs_ctx* ctx);   // captured steps bake kernel arguments in: drop them when those change

#define CS_CUDA(ctx, expr)                                                                       \
    do {                                                                                         \
        cudaError_t e__ = (expr);                                                                \
        if (e__ != cudaSuccess)                                                                  \
            return cs_fail(ctx, CS_E_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e__), \
                           __FILE__, __LINE__);                                                  \
    } while (0)

// topology (cs_topology.cu)
struct SeamInfo { int fb, eb, rev, code; };
SeamInfo cs_seam(int face, int edge);            // edge: 0 N, 1 S, 2 E, 3 W
void cs_resolve(int face, int64_t J, int64_t I, int N, int* fb, int64_t* Jb, int64_t* Ib, int* code);
int cs_build_halo(cs_ctx* ctx);
int cs_upload_metric(cs_ctx* ctx);               // (re)build the device tangent tables from ctx->xt
int cs_build_peer_plan(cs_ctx* ctx, int n_peers, const int32_t* tb, const int32_t* tc);

// kernel launchers (cs_halo.cu, cs_stage.cu, cs_diag.cu)
int cs_launch_halo(cs_ctx* ctx, void* h, void* u, void* v, int nfields_mask, cudaStream_t s);
int cs_launch_stage(cs_ctx* ctx, const void* const q0[3], const void* const qs[3], void* const qout[3],
                    double a, double b, double dt, int part, double* diag_out, int sums_in, cudaStream_t s);   // sums_in: diag_out gets
                                                                          // the sums of the stage input instead of its output
// one launch per step (cs_step_kernel.cu): whether cs_step may take it for these arguments, and the launch itself
bool cs_step_kernel_eligible(const cs_ctx* ctx, const void* h, const void* u, const void* v, const double* diag_out);
int cs_launch_step_kernel(cs_ctx* ctx, void* h, void* u, void* v, double dt, double* diag_out, cudaStream_t s);
int cs_launch_diag(cs_ctx* ctx, const void* h, const void* u, const void* v, double* out, cudaStream_t s);
int cs_launch_pack(cs_ctx* ctx, const void* h, const void* u, const void* v, int nfields, void* sendbuf,
                   cudaStream_t s);
int cs_launch_unpack(cs_ctx* ctx, const void* recvbuf, void* h, void* u, void* v, int nfields,
                     cudaStream_t s);

// Launch with the programmatic-stream-serialisation attribute: the kernel may be scheduled while its predecessor
// in the stream drains and must call cudaGridDependencySynchronize() before it touches anything the predecessor
// wrote (stage_march_kernel and halo_fill_kernel do; nothing else is launched this way).
template <typename... KArgs, typename... Args>
inline cudaError_t cs_launch_dependent(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, bool pdl,
                                       Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid, cfg.blockDim = block, cfg.dynamicSmemBytes = smem, cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr, cfg.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
