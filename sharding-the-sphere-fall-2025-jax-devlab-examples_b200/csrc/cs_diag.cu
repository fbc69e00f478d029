// K3 -- mass and total-energy reductions.
//
// Per-thread fp64 accumulation over a fixed grid-stride pattern, warp shuffles, one partial
// per block, and the last block to finish adds the partials in index order: the result is
// deterministic for a given configuration (no floating-point atomics anywhere).
#include "cs_internal.cuh"
#include "cs_cell.cuh"
#include "cs_reduce.cuh"

namespace {

constexpr int DIAG_THREADS = 256;
constexpr int DIAG_UNROLL = 4;

template <typename R>
__global__ void __launch_bounds__(DIAG_THREADS)
diag_kernel(const R* __restrict__ h, const R* __restrict__ u, const R* __restrict__ v, const R* __restrict__ topo,
            const double* __restrict__ xt, int n, int g, int m, int p, int T, const int32_t* __restrict__ tile_ids, int rows_per_block, double a2, double grav,
            double darea,
            double* __restrict__ partials, unsigned int* __restrict__ ticket, double* __restrict__ out) {
    // a block owns `rows_per_block` consecutive rows of one tile (no integer division per cell);
    // threads run along x, so loads are coalesced and every thread's summation order is fixed
    double mass = 0.0, energy = 0.0;
    const int blocks_per_tile = (n + rows_per_block - 1) / rows_per_block;
    for (int b = blockIdx.x; b < T * blocks_per_tile; b += gridDim.x) {
        const int t = b / blocks_per_tile;
        const int j0 = (b - t * blocks_per_tile) * rows_per_block;
        const int j1 = min(j0 + rows_per_block, n);
        const TileCoord tc = cs_tile_coord(tile_ids[t], p);
        const double* xt_x = xt + tc.tx * n + g;
        const double* xt_y = xt + tc.ty * n + g;
        for (int j = j0; j < j1; ++j) {
            const double Y = xt_y[j];
            const size_t row = (size_t)t * m * m + (size_t)(j + g) * m + g;
            // DIAG_UNROLL cells per thread and pass, all loads issued before the arithmetic: the kernel is bound by
            // memory latency (32 resident warps per SM, a chain of ~40 dependent fp64 operations per cell), so bytes
            // in flight per thread are what count.  Cells past the end of the row are loaded from a valid address
            // and zeroed: adding 0.0 leaves the sums bit-identical, and every thread's order stays fixed.
            for (int i0 = threadIdx.x; i0 < n; i0 += DIAG_UNROLL * DIAG_THREADS) {
                R hq[DIAG_UNROLL], uq[DIAG_UNROLL], vq[DIAG_UNROLL], tq[DIAG_UNROLL];
                double Xq[DIAG_UNROLL];
#pragma unroll
                for (int k = 0; k < DIAG_UNROLL; ++k) {
                    const int i = i0 + k * DIAG_THREADS;
                    const int ic = i < n ? i : i0;
                    const size_t o = row + ic;
                    hq[k] = h[o], uq[k] = u[o], vq[k] = v[o];
                    tq[k] = topo ? topo[o] : R(0);
                    Xq[k] = xt_x[ic];
                }
#pragma unroll
                for (int k = 0; k < DIAG_UNROLL; ++k) {
                    const bool in = i0 + k * DIAG_THREADS < n;
                    const Metric<double> M = cs_metric<double>(Xq[k], Y, a2);
                    const double hh = in ? (double)hq[k] : 0.0, uu = (double)uq[k], vv = (double)vq[k];
                    const double hs = (double)tq[k];
                    const Derived<double> D = cs_derive<double>(M, hh, uu, vv, hs, grav);
                    const double w = M.sg * darea;
                    mass += w * hh;
                    energy += w * (hh * D.K + 0.5 * grav * hh * hh + grav * hh * hs);
                }
            }
        }
    }
    cs_block_sum2<DIAG_THREADS>(mass, energy);
    cs_finish_partials<DIAG_THREADS>(partials, ticket, blockIdx.x, gridDim.x, mass, energy, out);
}

}  // namespace

int cs_launch_diag(cs_ctx* ctx, const void* h, const void* u, const void* v, double* out, cudaStream_t s) {
    // about 8 blocks per SM, each a whole number of rows of one tile
    const int64_t want = std::min<int64_t>(ctx->partial_cap, (int64_t)ctx->num_sms * 8);
    int rows_per_block = (int)(((int64_t)ctx->T * ctx->n + want - 1) / want);
    if (rows_per_block < 1) rows_per_block = 1;
    const int64_t blocks = std::min<int64_t>(want, (int64_t)ctx->T * ((ctx->n + rows_per_block - 1) / rows_per_block));
    const double a2 = ctx->radius * ctx->radius, darea = ctx->delta * ctx->delta;
    if (ctx->dtype == CS_F64)
        diag_kernel<double><<<(int)blocks, DIAG_THREADS, 0, s>>>((const double*)h, (const double*)u, (const double*)v,
                                                                ctx->has_topo ? (const double*)ctx->topo : nullptr, ctx->d_xt64,
                                                                ctx->n, ctx->g, ctx->m, ctx->p, ctx->T, ctx->d_tile_ids, rows_per_block, a2,
                                                                ctx->gravity, darea, ctx->d_partials + 4 * (size_t)ctx->partial_cap, ctx->d_ticket + 2, out);
    else
        diag_kernel<float><<<(int)blocks, DIAG_THREADS, 0, s>>>((const float*)h, (const float*)u, (const float*)v,
                                                               ctx->has_topo ? (const float*)ctx->topo : nullptr, ctx->d_xt64,
                                                               ctx->n, ctx->g, ctx->m, ctx->p, ctx->T, ctx->d_tile_ids, rows_per_block, a2,
                                                               ctx->gravity, darea, ctx->d_partials + 4 * (size_t)ctx->partial_cap, ctx->d_ticket + 2, out);
    ctx->launches++;
    CS_CUDA(ctx, cudaGetLastError());
    return CS_OK;
}

extern "C" int cs_diagnostics(cs_ctx* ctx, const void* h, const void* u, const void* v, double* out, void* stream) {
    if (!ctx || !h || !u || !v || !out) return cs_fail(ctx, CS_E_INVALID, "cs_diagnostics: NULL argument");
    CS_ON_DEVICE(ctx, ctx->device);
    return cs_launch_diag(ctx, h, u, v, out, (cudaStream_t)stream);
}
