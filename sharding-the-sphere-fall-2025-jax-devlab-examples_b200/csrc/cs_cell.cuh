// Per-cell arithmetic of the scheme, shared by every stage kernel
// variant and by the diagnostics kernel.  Host-callable too, so tests can run the very
// same expressions on the CPU against the NumPy oracle before any GPU time is spent.
#pragma once

#include <math.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define CS_HD __host__ __device__ __forceinline__
#else
#define CS_HD inline
#endif

#ifndef CS_LIM_NONE
#define CS_LIM_NONE 0
#define CS_LIM_MINMOD 1
#define CS_LIM_MC 2
#endif

template <typename R>
struct StageParams {
    // q0: state at the start of the step (point-wise), qs: stage input (stencil), out: stage output
    const R* h0; const R* u0; const R* v0;
    const R* hq; const R* uq; const R* vq;
    R* ho; R* uo; R* vo;
    const R* topo;            // bottom topography incl. ghosts, or nullptr
    const R* xt;              // tan(alpha) at padded face-level index (periodic continuation)
    const R* ainv;            // 1 / (1 + X^2)
    const int4* push;         // [T][4] ghost push descriptors, or nullptr
    double* partials;         // [work items][2] mass / energy partial sums of a stage with fused diagnostics, or nullptr
    unsigned int* ticket;     // ... its completion counter (zero between launches)
    double* diag_out;         // ... and where [mass, energy] go
    int sums_in;              // the sums are those of the stage INPUT (qs) instead of the output
    const int32_t* tile_ids;  // [T] tile id (face p^2 + ty p + tx) of every owned position, or nullptr: tile id = t0 + position
    int n, g, m, p, T, t0;
    int part;                 // 0 inner cells, 1 rim, 2 all
    int advect_only;
    R a, b, dt;               // out = a q0 + b (qs + dt L(qs)); a == 0 -> q0 is not read
    R rd, r2d;                // 1 / delta, 1 / (2 delta)
    R a2, ia2, grav, two_omega;
    R darea;                  // delta^2 (diagnostics weight)
    R cf[6][3];               // Coriolis coefficients per face
};

template <typename R> CS_HD R cs_rsqrt(R x);
template <> CS_HD float cs_rsqrt<float>(float x) {
#if defined(__CUDA_ARCH__)
    return rsqrtf(x);
#else
    return 1.0f / sqrtf(x);
#endif
}
template <> CS_HD double cs_rsqrt<double>(double x) {
#if defined(__CUDA_ARCH__)
    return rsqrt(x);
#else
    return 1.0 / sqrt(x);
#endif
}

CS_HD float cs_min(float a, float b) { return fminf(a, b); }
CS_HD double cs_min(double a, double b) { return a < b ? a : b; }   // no NaN handling needed: cheaper than fmin in fp64
CS_HD float cs_max(float a, float b) { return fmaxf(a, b); }
CS_HD double cs_max(double a, double b) { return a > b ? a : b; }
CS_HD float cs_abs(float a) { return fabsf(a); }
CS_HD double cs_abs(double a) { return fabs(a); }

// metric at a cell: everything the stencil needs from (X, Y)
template <typename R>
struct Metric {
    R X, Y, A, B, rho2, c, sg;   // c = a^2 A B / rho^4, sg = sqrt(G) = c rho
};

template <typename R>
CS_HD Metric<R> cs_metric(R X, R Y, R a2) {
    Metric<R> M;
    M.X = X, M.Y = Y;
    M.A = R(1) + X * X;
    M.B = R(1) + Y * Y;
    M.rho2 = M.A + Y * Y;
    const R rinv = cs_rsqrt<R>(M.rho2);
    const R r2inv = rinv * rinv;
    M.c = (a2 * M.A) * M.B * (r2inv * r2inv);
    M.sg = M.c * (M.rho2 * rinv);
    return M;
}

// derived cell quantities used by neighbours
template <typename R>
struct Derived {
    R m1, m2;     // sqrt(G) u^1, sqrt(G) u^2   (mass-flux velocities)
    R c1, c2;     // covariant velocity components
    R Bn;         // Bernoulli function g (h + hs) + K
    R K;          // kinetic energy per unit mass
};

template <typename R>
CS_HD Derived<R> cs_derive(const Metric<R>& M, R h, R u, R v, R hs, R grav) {
    Derived<R> D;
    const R g11 = M.c * M.A, g22 = M.c * M.B, g12 = -(M.c * M.X) * M.Y;
    D.m1 = M.sg * u;
    D.m2 = M.sg * v;
    D.c1 = g11 * u + g12 * v;
    D.c2 = g12 * u + g22 * v;
    D.K = R(0.5) * (u * D.c1 + v * D.c2);
    D.Bn = grav * (h + hs) + D.K;
    return D;
}

// Limited slope from the two one-sided differences.  Written with min / max only:
//   minmod(x, y, ...) = max(0, min(x, y, ...)) + min(0, max(x, y, ...))
// (the first term is non-zero only when all arguments are positive, the second only when all are
// negative), which gives the same value as sign * min(|.|) without compares and selects.
template <typename R, int LIM>
CS_HD R cs_slope(R dm, R dp) {
    if (LIM == CS_LIM_NONE) return R(0.5) * (dm + dp);
    if (LIM == CS_LIM_MINMOD) return cs_max(R(0), cs_min(dm, dp)) + cs_min(R(0), cs_max(dm, dp));
    const R a2 = dm + dm, b2 = dp + dp, c = R(0.5) * (dm + dp);
    return cs_max(R(0), cs_min(cs_min(a2, b2), c)) + cs_min(R(0), cs_max(cs_max(a2, b2), c));
}

// Upwind flux through the face between a left (L) and right (R) cell:
//   max(M, 0) hl + min(M, 0) hr  ==  (M (hl + hr) + |M| (hl - hr)) / 2,   M = (mL + mR) / 2.
// The second form needs no min / max (|.| is a free operand modifier) and is exactly antisymmetric
// under exchange of the two sides, which keeps seam fluxes single valued.
template <typename R>
CS_HD R cs_face_flux(R mL, R mR, R hL, R sL, R hR, R sR) {
    const R Fm = R(0.25) * (mL + mR);
    const R left = hL + R(0.5) * sL, right = hR - R(0.5) * sR;
    return Fm * (left + right) + cs_abs(Fm) * (left - right);
}

struct TileCoord { int face, ty, tx; };
CS_HD TileCoord cs_tile_coord(int tile_global, int p) {
    TileCoord c;
    c.face = tile_global / (p * p);
    const int r = tile_global - c.face * p * p;
    c.ty = r / p;
    c.tx = r - c.ty * p;
    return c;
}

// Tile id of shard position t: a table look-up only when the tiles are not in plain order (the look-up is a dependent
// global load at the very start of every warp: 1.4 us per fp32 step at C1536).
template <typename R> CS_HD int cs_tile_id(const StageParams<R>& P, int t) { return P.tile_ids ? P.tile_ids[t] : t + P.t0; }

// What a stage cell hands back to a caller that also wants the conservation sums: the new values, the
// topography and the cell's tangents.
template <typename R>
struct CellOut {
    R h, u, v, hs, X, Y;      // new values, topography, tangents
    R h_in, u_in, v_in;       // the stage input at the same cell
};

// One interior cell (j, i) of local tile t, everything from global memory.  This is the
// reference formulation of the stage on the GPU (CS_KERNEL_NAIVE); the fast kernels compute
// the same expressions with the loads and the derived quantities shared between cells.
template <typename R, int LIM>
CS_HD void cs_stage_cell(const StageParams<R>& P, int t, int j, int i, CellOut<R>* out = nullptr) {
    const int m = P.m, g = P.g;
    const TileCoord tc = cs_tile_coord(cs_tile_id(P, t), P.p);
    const R* xt = P.xt + tc.tx * P.n;
    const R* yt = P.xt + tc.ty * P.n;
    const int jt = j + g, it = i + g;
    const size_t base = (size_t)t * m * m;
    const size_t o = base + (size_t)jt * m + it;
    const R* hq = P.hq;

    const R hC = hq[o];
    const R hW1 = hq[o - 1], hW2 = hq[o - 2], hE1 = hq[o + 1], hE2 = hq[o + 2];
    const R hS1 = hq[o - m], hS2 = hq[o - 2 * m], hN1 = hq[o + m], hN2 = hq[o + 2 * m];

    const Metric<R> MC = cs_metric<R>(xt[it], yt[jt], P.a2);
    const R uC = P.uq[o], vC = P.vq[o];
    const Derived<R> DC = cs_derive<R>(MC, hC, uC, vC, P.topo ? P.topo[o] : R(0), P.grav);
    const Derived<R> DW = cs_derive<R>(cs_metric<R>(xt[it - 1], yt[jt], P.a2), hW1, P.uq[o - 1], P.vq[o - 1],
                                       P.topo ? P.topo[o - 1] : R(0), P.grav);
    const Derived<R> DE = cs_derive<R>(cs_metric<R>(xt[it + 1], yt[jt], P.a2), hE1, P.uq[o + 1], P.vq[o + 1],
                                       P.topo ? P.topo[o + 1] : R(0), P.grav);
    const Derived<R> DS = cs_derive<R>(cs_metric<R>(xt[it], yt[jt - 1], P.a2), hS1, P.uq[o - m], P.vq[o - m],
                                       P.topo ? P.topo[o - m] : R(0), P.grav);
    const Derived<R> DN = cs_derive<R>(cs_metric<R>(xt[it], yt[jt + 1], P.a2), hN1, P.uq[o + m], P.vq[o + m],
                                       P.topo ? P.topo[o + m] : R(0), P.grav);

    const R sxW = cs_slope<R, LIM>(hW1 - hW2, hC - hW1);
    const R sxC = cs_slope<R, LIM>(hC - hW1, hE1 - hC);
    const R sxE = cs_slope<R, LIM>(hE1 - hC, hE2 - hE1);
    const R syS = cs_slope<R, LIM>(hS1 - hS2, hC - hS1);
    const R syC = cs_slope<R, LIM>(hC - hS1, hN1 - hC);
    const R syN = cs_slope<R, LIM>(hN1 - hC, hN2 - hN1);

    const R FE = cs_face_flux<R>(DC.m1, DE.m1, hC, sxC, hE1, sxE);
    const R FW = cs_face_flux<R>(DW.m1, DC.m1, hW1, sxW, hC, sxC);
    const R GN = cs_face_flux<R>(DC.m2, DN.m2, hC, syC, hN1, syN);
    const R GS = cs_face_flux<R>(DS.m2, DC.m2, hS1, syS, hC, syC);

    const R isg = R(1) / MC.sg;
    const R dh = -(((FE - FW) + (GN - GS)) * P.rd) * isg;

    R du = R(0), dv = R(0);
    if (!P.advect_only) {
        const R fsg = P.two_omega * (P.cf[tc.face][0] + P.cf[tc.face][1] * MC.X + P.cf[tc.face][2] * MC.Y) * MC.c;
        const R eta = ((DE.c2 - DW.c2) - (DN.c1 - DS.c1)) * P.r2d + fsg;
        const R T1 = eta * vC - (DE.Bn - DW.Bn) * P.r2d;
        const R T2 = -eta * uC - (DN.Bn - DS.Bn) * P.r2d;
        const R ax = P.ainv[tc.tx * P.n + it], ay = P.ainv[tc.ty * P.n + jt];
        const R q = MC.rho2 * P.ia2;
        const R gi11 = q * ax, gi22 = q * ay, gi12 = (q * (MC.X * MC.Y)) * (ax * ay);
        du = gi11 * T1 + gi12 * T2;
        dv = gi12 * T1 + gi22 * T2;
    }
    R h1 = P.b * (hC + P.dt * dh), u1 = P.b * (uC + P.dt * du), v1 = P.b * (vC + P.dt * dv);
    if (P.a != R(0)) {
        h1 += P.a * P.h0[o];
        u1 += P.a * P.u0[o];
        v1 += P.a * P.v0[o];
    }
    P.ho[o] = h1;
    P.uo[o] = u1;
    P.vo[o] = v1;
    if (out) {
        out->h = h1, out->u = u1, out->v = v1;
        out->h_in = hC, out->u_in = uC, out->v_in = vC;
        out->hs = P.topo ? P.topo[o] : R(0);
        out->X = MC.X, out->Y = MC.Y;
    }
}

// Contribution of one cell to (mass, energy), without the factor delta^2: the all-fp64 per-cell form the
// reduction kernel uses.
CS_HD void cs_cell_sums(double X, double Y, double a2, double grav, double h, double u, double v, double hs, double& mass,
                        double& energy) {
    const Metric<double> M = cs_metric<double>(X, Y, a2);
    const Derived<double> D = cs_derive<double>(M, h, u, v, hs, grav);
    mass += M.sg * h;
    energy += M.sg * (h * D.K + 0.5 * grav * h * h + grav * h * hs);
}
