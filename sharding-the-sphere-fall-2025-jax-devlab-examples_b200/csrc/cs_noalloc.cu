// cs_diagnostics_noalloc -- mass / energy reduction for callers that must not allocate.
//
// Who this is for.  cs_diagnostics() needs a cs_ctx, and creating a context calls cudaMalloc
// (metric tables, gather tables, partial-sum scratch).  Some users of this library cannot do
// that: a handler that runs inside a framework which has already reserved the GPU's memory for
// its own allocator (JAX/XLA preallocates most of the device, so a plugin's cudaMalloc can fail
// with out-of-memory even though the framework has plenty free), and monitoring hooks on a
// latency-critical thread, where cudaMalloc/cudaFree take a device-wide lock and may wait for
// all running work.  Such a caller already holds h, u, v; this entry point lets it get the two
// conservation numbers with no context and no allocation at all.
//
// How.  The only scratch the reduction needs -- one partial (mass, energy) pair per thread block
// and a ticket counter -- lives in a file-scope static __device__ array of fixed size below.
// Every GPU that loads the library carries that array (about 128 KB) until the process exits.
// The tangent table a context would hold is replaced by evaluating tan() per cell in the kernel.
//
// Limits and rules (also in include/cubesphere.h):
//   * at most CS_NOALLOC_MAX_CELLS interior cells per call; larger shards are rejected;
//   * the metric is the default equiangular one (tan() of the cell-centre angles); a context's custom tangent
//     table is not known here and not honoured;
//   * the calling thread's current device is restored on return;
//   * the first call on each device looks up the array's device address
//     (cudaGetSymbolAddress, with that device current) and remembers it; make that first call
//     before starting any stream capture on that device;
//   * all calls on a device share the one array: it still holds the previous call's partials, the
//     ticket is cleared at the start of every call, and two calls whose kernels overlap on the GPU
//     (different streams, or different host threads) race on it -- serialise them.
#include <mutex>

#include "cs_internal.cuh"
#include "cs_cell.cuh"

#define CS_NOALLOC_THREADS 256
#define CS_NOALLOC_CELLS_PER_THREAD 16
#define CS_NOALLOC_MAX_BLOCKS 8192
#define CS_NOALLOC_MAX_CELLS ((long long)CS_NOALLOC_MAX_BLOCKS * CS_NOALLOC_THREADS * CS_NOALLOC_CELLS_PER_THREAD)
#define CS_NOALLOC_MAX_DEVICES 64

// [2 * MAX_BLOCKS] partial sums, then one 8-byte slot whose first 4 bytes are the ticket counter.
// In host code this name is only a handle for cudaGetSymbolAddress, never a usable pointer.
static __device__ __align__(256) double cs_noalloc_scratch[2 * CS_NOALLOC_MAX_BLOCKS + 1];

namespace {

std::mutex g_setup_mutex;
double* g_scratch_ptr[CS_NOALLOC_MAX_DEVICES];
bool g_scratch_ready[CS_NOALLOC_MAX_DEVICES];

__device__ __forceinline__ double na_warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    return v;
}

__device__ __forceinline__ void na_block_sum2(double& a, double& b) {
    __shared__ double sa[CS_NOALLOC_THREADS / 32], sb[CS_NOALLOC_THREADS / 32];
    a = na_warp_sum(a);
    b = na_warp_sum(b);
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    if (l == 0) sa[w] = a, sb[w] = b;
    __syncthreads();
    if (w == 0) {
        a = l < CS_NOALLOC_THREADS / 32 ? sa[l] : 0.0;
        b = l < CS_NOALLOC_THREADS / 32 ? sb[l] : 0.0;
        a = na_warp_sum(a);
        b = na_warp_sum(b);
    }
    __syncthreads();
}

// tan(alpha) at face-level interior index i, built like the context's table: the upper half from
// tan(), the lower half by exact antisymmetry.
__device__ __forceinline__ double na_tangent(int i, int N) {
    const double pi = 3.14159265358979323846;
    const int k = i >= N - 1 - i ? i : N - 1 - i;
    if ((N & 1) && i == N / 2) return 0.0;
    const double t = tan(-0.25 * pi + (k + 0.5) * (0.5 * pi / N));
    return i == k ? t : -t;
}

template <typename R>
__global__ void __launch_bounds__(CS_NOALLOC_THREADS)
diag_noalloc_kernel(const R* __restrict__ h, const R* __restrict__ u, const R* __restrict__ v, const R* __restrict__ topo,
                    int N, int n, int g, int m, int p, int T, int t0, double a2, double grav, double darea,
                    double* __restrict__ scratch, double* __restrict__ out) {
    double* partials = scratch;
    unsigned int* ticket = reinterpret_cast<unsigned int*>(scratch + 2 * CS_NOALLOC_MAX_BLOCKS);
    double mass = 0.0, energy = 0.0;
    const long long cells = (long long)T * n * n;
    const long long base = (long long)blockIdx.x * (CS_NOALLOC_THREADS * CS_NOALLOC_CELLS_PER_THREAD);
#pragma unroll 1
    for (int k = 0; k < CS_NOALLOC_CELLS_PER_THREAD; ++k) {
        const long long c = base + (long long)k * CS_NOALLOC_THREADS + threadIdx.x;
        if (c >= cells) break;
        const int t = (int)(c / ((long long)n * n));
        const int r = (int)(c - (long long)t * n * n);
        const int j = r / n, i = r - j * n;
        const TileCoord tc = cs_tile_coord(t + t0, p);
        const size_t o = (size_t)t * m * m + (size_t)(j + g) * m + (i + g);
        const Metric<double> M = cs_metric<double>(na_tangent(tc.tx * n + i, N), na_tangent(tc.ty * n + j, N), a2);
        const double hh = (double)h[o], uu = (double)u[o], vv = (double)v[o];
        const double hs = topo ? (double)topo[o] : 0.0;
        const Derived<double> D = cs_derive<double>(M, hh, uu, vv, hs, grav);
        const double w = M.sg * darea;
        mass += w * hh;
        energy += w * (hh * D.K + 0.5 * grav * hh * hh + grav * hh * hs);
    }
    na_block_sum2(mass, energy);
    __shared__ bool last;
    if (threadIdx.x == 0) {
        partials[2 * blockIdx.x] = mass;
        partials[2 * blockIdx.x + 1] = energy;
        __threadfence();
        last = atomicAdd(ticket, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (last) {   // the last block to arrive adds the partials in index order: deterministic
        __threadfence();
        double a = 0.0, b = 0.0;
        for (int k = threadIdx.x; k < (int)gridDim.x; k += CS_NOALLOC_THREADS) {
            a += __ldcg(partials + 2 * k);                // written by other blocks: read through L2, not a stale L1 line
            b += __ldcg(partials + 2 * k + 1);
        }
        na_block_sum2(a, b);
        if (threadIdx.x == 0) out[0] = a, out[1] = b;
    }
}

// Once per device, with that device current: turn the symbol into a device address.
int noalloc_setup(int device, double** ptr) {
    std::lock_guard<std::mutex> lock(g_setup_mutex);
    if (!g_scratch_ready[device]) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, cs_noalloc_scratch);
        if (e != cudaSuccess || !p)
            return cs_fail(nullptr, CS_E_CUDA, "cs_diagnostics_noalloc: cudaGetSymbolAddress failed on device %d: %s", device,
                           cudaGetErrorString(e));
        g_scratch_ptr[device] = static_cast<double*>(p);
        g_scratch_ready[device] = true;
    }
    *ptr = g_scratch_ptr[device];
    return CS_OK;
}

}  // namespace

extern "C" int cs_diagnostics_noalloc(int dtype, int N, int tiles_per_edge, int halo, int tile_begin, int tile_count, double radius,
                                      double gravity, const void* h, const void* u, const void* v, const void* hs, double* out,
                                      int device, void* stream) {
    if (!h || !u || !v || !out) return cs_fail(nullptr, CS_E_INVALID, "cs_diagnostics_noalloc: NULL argument");
    if (dtype != CS_F32 && dtype != CS_F64) return cs_fail(nullptr, CS_E_INVALID, "cs_diagnostics_noalloc: unknown dtype %d", dtype);
    const int p = tiles_per_edge;
    if (p < 1 || N < 2 || N % p) return cs_fail(nullptr, CS_E_INVALID, "cs_diagnostics_noalloc: N = %d must be >= 2 and divisible by tiles_per_edge = %d", N, p);
    const int n = N / p, T_all = 6 * p * p, T = tile_count == 0 ? T_all : tile_count;
    if (halo < 1 || halo > CS_MAX_HALO) return cs_fail(nullptr, CS_E_INVALID, "cs_diagnostics_noalloc: halo = %d must be in [1, %d]", halo, CS_MAX_HALO);
    if (tile_begin < 0 || T < 1 || tile_begin + T > T_all)
        return cs_fail(nullptr, CS_E_INVALID, "cs_diagnostics_noalloc: tile range [%d, %d) is outside [0, %d)", tile_begin, tile_begin + T, T_all);
    if (device < 0 || device >= CS_NOALLOC_MAX_DEVICES) return cs_fail(nullptr, CS_E_INVALID, "cs_diagnostics_noalloc: device %d out of range", device);
    const long long cells = (long long)T * n * n;
    if (cells > CS_NOALLOC_MAX_CELLS)
        return cs_fail(nullptr, CS_E_INVALID,
                       "cs_diagnostics_noalloc: %lld interior cells exceed the fixed scratch (%lld cells = %d block partials); "
                       "use cs_diagnostics with a context, or reduce per shard",
                       cells, CS_NOALLOC_MAX_CELLS, CS_NOALLOC_MAX_BLOCKS);
    CS_ON_DEVICE(nullptr, device);                       // the caller's current device is put back on return
    cudaError_t e;
    double* scratch = nullptr;
    int rc = noalloc_setup(device, &scratch);
    if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    // the array keeps whatever the previous call left: only the ticket has to start at zero
    e = cudaMemsetAsync(scratch + 2 * CS_NOALLOC_MAX_BLOCKS, 0, sizeof(double), s);
    if (e != cudaSuccess) return cs_fail(nullptr, CS_E_CUDA, "cs_diagnostics_noalloc: clearing the ticket failed: %s", cudaGetErrorString(e));
    const int per_block = CS_NOALLOC_THREADS * CS_NOALLOC_CELLS_PER_THREAD;
    const int blocks = (int)((cells + per_block - 1) / per_block);
    const int m = n + 2 * halo;
    const double a = radius != 0 ? radius : 6.37122e6, grav = gravity != 0 ? gravity : 9.80616;
    const double delta = 0.5 * 3.14159265358979323846 / N;
    if (dtype == CS_F64)
        diag_noalloc_kernel<double><<<blocks, CS_NOALLOC_THREADS, 0, s>>>((const double*)h, (const double*)u, (const double*)v,
                                                                         (const double*)hs, N, n, halo, m, p, T, tile_begin, a * a,
                                                                         grav, delta * delta, scratch, out);
    else
        diag_noalloc_kernel<float><<<blocks, CS_NOALLOC_THREADS, 0, s>>>((const float*)h, (const float*)u, (const float*)v,
                                                                        (const float*)hs, N, n, halo, m, p, T, tile_begin, a * a, grav,
                                                                        delta * delta, scratch, out);
    e = cudaGetLastError();
    if (e != cudaSuccess) return cs_fail(nullptr, CS_E_CUDA, "cs_diagnostics_noalloc: launch failed: %s", cudaGetErrorString(e));
    return CS_OK;
}
