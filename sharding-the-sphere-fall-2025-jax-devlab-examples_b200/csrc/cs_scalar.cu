// Scalar demo workload on the halo machinery: explicit diffusion of a tracer T on the cubed sphere.
//
// The reference's "proof that sharding works" is a temperature field diffusing away from a checkerboard heat source on
// the top panel and crossing the four seams of that panel smoothly (deck slides 12 and 17).  This is that workload as
// a regression case for the scalar ghost fill: a 5-point flux-form stencil
//     T' = T + kdt / (sqrt(G) delta^2) [ d_a( W1 d_a T ) + d_b( W2 d_b T ) ] + dts S,
//     W1 = sqrt(G) g^11, W2 = sqrt(G) g^22 averaged to the cell faces,
// i.e. the Laplace-Beltrami operator without its cross-derivative (g^12) terms, which would need the corner ghosts no
// panel owns.  With the periodic metric continuation every face weight and every difference across a seam is the same
// number on both sides, so each seam flux is single valued and the total heat  sum sqrt(G) T  is conserved to
// round-off; what is lost by dropping g^12 is accuracy towards the cube corners, not conservation.
// One thread per cell, straight from global memory: the field is one word per cell and the stencil is five loads.
#include "cs_internal.cuh"
#include "cs_cell.cuh"

namespace {

template <typename R>
__global__ void __launch_bounds__(256)
diffuse_kernel(const R* __restrict__ T, R* __restrict__ out, const R* __restrict__ src, const double* __restrict__ xt, int n, int g,
               int m, int p, const int32_t* __restrict__ tile_ids, double a2, double kdt_over_d2, double dts) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y * blockDim.y + threadIdx.y;
    const int t = blockIdx.z;
    if (i >= n || j >= n) return;
    const TileCoord tc = cs_tile_coord(tile_ids[t], p);
    const double* xx = xt + tc.tx * n + g + i;            // tangents at columns i-1, i, i+1 (ghosts: periodic table)
    const double* yy = xt + tc.ty * n + g + j;
    const size_t o = (size_t)t * m * m + (size_t)(j + g) * m + (i + g);
    // sqrt(G) g^11 = rho B / ... evaluated from the metric: sg * rho2 / (a^2 A), sg * rho2 / (a^2 B)
    auto w1 = [&](double X, double Y) { const Metric<double> M = cs_metric<double>(X, Y, a2); return M.sg * M.rho2 / (a2 * M.A); };
    auto w2 = [&](double X, double Y) { const Metric<double> M = cs_metric<double>(X, Y, a2); return M.sg * M.rho2 / (a2 * M.B); };
    const double X = xx[0], Y = yy[0];
    const double c = (double)T[o];
    const double wC1 = w1(X, Y), wC2 = w2(X, Y);
    const double FE = 0.5 * (wC1 + w1(xx[1], Y)) * ((double)T[o + 1] - c);
    const double FW = 0.5 * (w1(xx[-1], Y) + wC1) * (c - (double)T[o - 1]);
    const double GN = 0.5 * (wC2 + w2(X, yy[1])) * ((double)T[o + m] - c);
    const double GS = 0.5 * (w2(X, yy[-1]) + wC2) * (c - (double)T[o - m]);
    const double sg = cs_metric<double>(X, Y, a2).sg;
    double r = c + kdt_over_d2 * (((FE - FW) + (GN - GS)) / sg);
    if (src) r += dts * (double)src[o];
    out[o] = (R)r;
}

}  // namespace

extern "C" int cs_scalar_diffuse(cs_ctx* ctx, const void* T, void* Tout, const void* source, double kappa_dt, double dt_source,
                                 void* stream) {
    if (!ctx || !T || !Tout) return cs_fail(ctx, CS_E_INVALID, "cs_scalar_diffuse: NULL argument");
    if (T == Tout) return cs_fail(ctx, CS_E_INVALID, "cs_scalar_diffuse: Tout must not alias T (stencil input)");
    CS_ON_DEVICE(ctx, ctx->device);
    cudaStream_t s = (cudaStream_t)stream;
    const dim3 block(32, 8), grid((ctx->n + 31) / 32, (ctx->n + 7) / 8, ctx->T);
    const double a2 = ctx->radius * ctx->radius, k = kappa_dt / (ctx->delta * ctx->delta);
    if (ctx->dtype == CS_F64)
        diffuse_kernel<double><<<grid, block, 0, s>>>((const double*)T, (double*)Tout, (const double*)source, ctx->d_xt64, ctx->n, ctx->g,
                                                     ctx->m, ctx->p, ctx->d_tile_ids, a2, k, dt_source);
    else
        diffuse_kernel<float><<<grid, block, 0, s>>>((const float*)T, (float*)Tout, (const float*)source, ctx->d_xt64, ctx->n, ctx->g,
                                                    ctx->m, ctx->p, ctx->d_tile_ids, a2, k, dt_source);
    ctx->launches++;
    CS_CUDA(ctx, cudaGetLastError());
    return cs_launch_halo(ctx, Tout, nullptr, nullptr, 1, s);      // local ghosts; a sharded caller exchanges the rest
}
