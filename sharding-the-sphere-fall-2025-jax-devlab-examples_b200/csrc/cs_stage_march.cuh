// CS_KERNEL_MARCH -- register-marching fused RK stage.
//
// A warp owns a strip of 30 column groups (V adjacent cells per thread, one 8/16-byte vector
// access per variable per row) and marches along y.  Everything a cell shares with its neighbours
// is computed ONCE by the owning thread and kept in registers:
//   * y-neighbours come from the sliding window of rows (J-1, J, J+1, J+2) held by the same thread,
//     the north-face flux of row J becomes the south-face flux of row J+1;
//   * x-neighbours come from the adjacent lanes by warp shuffles (2 h values, m1, u_2 and the
//     Bernoulli function per side); lanes 0 and 31 are halo lanes that compute but do not store.
// No shared memory and no block-level synchronisation; loads for row J+3 are issued before the
// arithmetic of row J, so each warp keeps about two rows of traffic in flight.
// HBM traffic per cell is the algorithmic 2V / 3V words plus the strip overlap (4 extra rows per
// `rows`, 2 of 32 lanes), most of which hits L2.
#pragma once

#include <type_traits>

#include "cs_cell.cuh"

struct MarchGeom {
    int prefetch;             // L2 prefetch distance in rows (0 = none)
    int xlo, xhi, ylo, yhi;   // interior index ranges [lo, hi) to update
    int rows;                 // rows per strip
    int grp0;                 // first column group that contains an output column
    int ngroups;              // column groups per row (m / V)
    int nchunks;              // warps along x
    int nstrips;              // strips along y
    int nitems;               // nchunks * nstrips * tiles (one warp each)
};

__device__ __forceinline__ void march_prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

template <typename R, int V> struct VecIO;
template <> struct VecIO<float, 2> {
    static __device__ __forceinline__ void ld(const float* p, float (&o)[2]) {
        const float2 t = *reinterpret_cast<const float2*>(p);
        o[0] = t.x, o[1] = t.y;
    }
    static __device__ __forceinline__ void st(float* p, const float (&o)[2]) {
        *reinterpret_cast<float2*>(p) = make_float2(o[0], o[1]);
    }
};
template <> struct VecIO<float, 4> {
    static __device__ __forceinline__ void ld(const float* p, float (&o)[4]) {
        const float4 t = *reinterpret_cast<const float4*>(p);
        o[0] = t.x, o[1] = t.y, o[2] = t.z, o[3] = t.w;
    }
    static __device__ __forceinline__ void st(float* p, const float (&o)[4]) {
        *reinterpret_cast<float4*>(p) = make_float4(o[0], o[1], o[2], o[3]);
    }
};
template <> struct VecIO<double, 2> {
    static __device__ __forceinline__ void ld(const double* p, double (&o)[2]) {
        const double2 t = *reinterpret_cast<const double2*>(p);
        o[0] = t.x, o[1] = t.y;
    }
    static __device__ __forceinline__ void st(double* p, const double (&o)[2]) {
        *reinterpret_cast<double2*>(p) = make_double2(o[0], o[1]);
    }
};

template <> struct VecIO<double, 1> {
    static __device__ __forceinline__ void ld(const double* p, double (&o)[1]) { o[0] = *p; }
    static __device__ __forceinline__ void st(double* p, const double (&o)[1]) { *p = o[0]; }
};

#define CS_MARCH_THREADS 128
#define CS_MARCH_OUT_LANES 30

// derived quantities of one row segment (V cells) from raw values
template <typename R, int V>
struct RowDerived {
    R m1[V], m2[V], c1[V], c2[V], Bn[V];
    R isg[V], q[V], fsg[V];   // 1/sqrt(G), rho^2/a^2, f sqrt(G): only used when the row is the output row
    R sg[V], cc[V];           // sqrt(G) and a^2 A B / rho^4: read by the fused diagnostics only (dead otherwise)
};

// Half of the limited slope (what the face states need): scaling by 2 is exact, so
// minmod(dm, dp, (dm + dp) / 4) is bit-identical to cs_slope(dm, dp) / 2.
template <typename R, int LIM>
__device__ __forceinline__ R march_half_slope(R dm, R dp) {
    // max(0, lo) + min(0, hi) with lo <= hi is the median of (lo, 0, hi) = max(lo, min(hi, 0)): bit-identical,
    // one addition fewer
    if (LIM == CS_LIM_NONE) return R(0.25) * (dm + dp);
    if (LIM == CS_LIM_MINMOD) return R(0.5) * cs_max(cs_min(dm, dp), cs_min(cs_max(dm, dp), R(0)));
    const R c = R(0.25) * (dm + dp);
    return cs_max(cs_min(cs_min(dm, dp), c), cs_min(cs_max(cs_max(dm, dp), c), R(0)));
}
// fp64 has no min / max instruction (each is a compare and two selects), so the double version takes the
// minimum of the magnitudes and restores the common sign: same values, two thirds of the instructions.
template <>
__device__ __forceinline__ double march_half_slope<double, CS_LIM_MC>(double dm, double dp) {
    const double c = 0.25 * (dm + dp);
    const double mag = cs_min(cs_min(fabs(dm), fabs(dp)), fabs(c));
    return dm * dp > 0.0 ? copysign(mag, dm) : 0.0;
}
template <>
__device__ __forceinline__ double march_half_slope<double, CS_LIM_MINMOD>(double dm, double dp) {
    const double mag = 0.5 * cs_min(fabs(dm), fabs(dp));
    return dm * dp > 0.0 ? copysign(mag, dm) : 0.0;
}

// Twice the upwind face flux (the factor 1/2 is folded into MarchConst::kisg); hsL, hsR are half slopes, M the SUM of
// the two cells' mass-flux velocities.  M (hL + hsL) for M > 0, M (hR - hsR) otherwise: what the oracle's
// Fm (l + r) + |Fm| (l - r), Fm = M / 4, evaluates to up to one rounding, with the compare and the select on the ALU
// pipe instead of two more FMA-pipe operations.  (Scaling by powers of two is exact: carrying 2 F instead of 4 F saves
// an addition per face and changes no bit of the result.)
template <typename R>
__device__ __forceinline__ R march_flux4(R mL, R mR, R hL, R hsL, R hR, R hsR) {
    const R M = mL + mR, left = hL + hsL, right = hR - hsR;
    return M * (M > R(0) ? left : right);
}

// Constants folded once per thread so that the row loop multiplies as little as possible:
//   kq   = b dt / (2 delta a^2)   (momentum: out = a q0 + b u + kq rho^2 (g^ij-shape) T'),
//   kisg = b dt / (delta a^2)     (continuity: out = a q0 + b h - isg' (dF + dG)),
//   kf   = 2 Omega 2 delta        (f sqrt(G) in units of the undivided differences).
template <typename R, int V>
struct MarchConst {
    R a2A[V], cfx[V], Xax[V], ax[V], X[V];
    R kq, kisg, kf2;   // kf2 = kf * cf[2]
};

// Conservation sums of the stage INPUT (SUMS == 2), taken where the row is derived: sqrt(G) h and
// sqrt(G) h (Bn - g h / 2) = sqrt(G) h (K + g h / 2 + g hs) are two multiplications and one FMA away from values the
// derivation has in hand anyway, and nothing has to be kept for later.  `count` says whether this row belongs to the
// warp's own output rows (every row is derived by two strips, and by none for its ghost rows); om masks the columns.
struct NoSums {};
template <typename R, int V>
struct InSums {
    const bool (&om)[V];
    bool count;
    double& mass;
    double& energy;
};

template <typename R, int V, typename SUMS = NoSums>
__device__ __forceinline__ void march_derive(const StageParams<R>& P, const MarchConst<R, V>& K, R Y, R ay,
                                             const R (&h)[V], const R (&u)[V], const R (&v)[V], const R (&hs)[V],
                                             RowDerived<R, V>& D, const SUMS& sums = SUMS()) {
    const R Y2 = Y * Y, B = R(1) + Y2, ayk = ay * K.kisg, cfy = K.kf2 * Y;
    R pm = R(0), pe = R(0);
#pragma unroll
    for (int c = 0; c < V; ++c) {
        const R rho2 = K.X[c] * K.X[c] + B;
        const R rinv = cs_rsqrt<R>(rho2);
        const R r2 = rinv * rinv;
        const R cc = (K.a2A[c] * B) * (r2 * r2);
        const R rho = rho2 * rinv;
        const R sg = cc * rho;
        D.m1[c] = sg * u[c];
        D.m2[c] = sg * v[c];
        const R XY = K.X[c] * Y;
        D.c1[c] = cc * ((rho2 - Y2) * u[c] - XY * v[c]);
        D.c2[c] = cc * (B * v[c] - XY * u[c]);
        const R Kin = R(0.5) * (u[c] * D.c1[c] + v[c] * D.c2[c]);
        D.Bn[c] = P.grav * (h[c] + hs[c]) + Kin;
        D.q[c] = rho2 * K.kq;
        D.isg[c] = (rho2 * rho) * (K.ax[c] * ayk);          // b dt / delta * rho^3 / (a^2 A B)
        D.fsg[c] = (K.cfx[c] + cfy) * cc;
        D.sg[c] = sg;
        D.cc[c] = cc;
        if constexpr (!std::is_same<SUMS, NoSums>::value) {
            const R w = sg * h[c];
            if (sums.om[c]) pm += w, pe += w * (D.Bn[c] - (R(0.5) * P.grav) * h[c]);
        }
    }
    if constexpr (!std::is_same<SUMS, NoSums>::value) {
        // the row's V cells are added in the state's precision, the running sums are kept in fp64
        if (sums.count) sums.mass += (double)pm, sums.energy += (double)pe;
    }
}

// Fused conservation sums (K3 inside the last stage of a step): mass sqrt(G) h and energy
// sqrt(G) h (K + g h / 2 + g hs) of the V freshly computed cells of one row, from values the thread already
// holds.  Returned without the constant factors (delta^2, and 1/2 for the energy), which the caller applies once
// per thread.  2 K = u_i u^i = cc (A u^2 - 2 X Y u v + B v^2).  Arithmetic in the state's precision per cell, the
// row's V cells are added in that precision and the running sums are kept in fp64.
template <typename R, int V>
__device__ __forceinline__ void march_row_diag(const StageParams<R>& P, const MarchConst<R, V>& K, const RowDerived<R, V>& D,
                                               R Y, const R (&h)[V], const R (&u)[V], const R (&v)[V], const R (&hs)[V],
                                               const bool (&om)[V], double& mass, double& energy2) {
    const R Bq = R(1) + Y * Y, a2B = P.a2 * Bq, a2Y2 = (P.a2 + P.a2) * Y;
    R pm = R(0), pe = R(0);
#pragma unroll
    for (int c = 0; c < V; ++c) {
        const R quad = K.a2A[c] * (u[c] * u[c]) - (a2Y2 * K.X[c]) * (u[c] * v[c]) + a2B * (v[c] * v[c]);
        // metric of the output row: kept from its derivation one iteration ago (fp32), or formed again (fp64: eight more
        // live double registers in a 255-register kernel cost more than the rsqrt -- 219 vs 236 us per launch at C1536;
        // in fp32 it is the other way round, 112 vs 116)
        R ccq, sgq;
        if constexpr (sizeof(R) == 8) {
            const R rho2 = K.X[c] * K.X[c] + Bq;
            const R rinv = cs_rsqrt<R>(rho2), r2 = rinv * rinv;
            ccq = (K.a2A[c] * Bq) * (r2 * r2), sgq = ccq * (rho2 * rinv);
        } else {
            ccq = D.cc[c], sgq = D.sg[c];
        }
        const R k2 = (ccq * P.ia2) * quad;                                       // 2 K
        const R col = k2 + P.grav * (h[c] + (hs[c] + hs[c]));                    // 2 (K + g h / 2 + g hs)
        const R w = sgq * h[c];
        pm += om[c] ? w : R(0);                                                  // halo lanes and ghost columns do not count
        pe += om[c] ? w * col : R(0);
    }
    mass += (double)pm;
    energy2 += (double)pe;
}

// Fused halo: a freshly computed boundary cell is also stored into the ghost cell of the neighbouring
// tile that mirrors it (same index mapping and signed permutation as the halo kernel, read from the
// per-(tile, edge) descriptor), so no separate ghost-fill launch is needed after the stage.
// A PushMap is that mapping for one (edge, depth): the ghost of the cell at along-edge index k lives at
// element offset base + k * stride.  e: 0 N, 1 S, 2 E, 3 W.
struct PushMap {
    unsigned base;
    int stride;
    int code;      // -1: the neighbour is not on this device (no push)
};

template <typename R>
__device__ __forceinline__ PushMap march_push_map(const StageParams<R>& P, int t, int e, int d) {
    const int4 pd = P.push[t * 4 + e];                    // (dst tile or -1, dst edge, reverse, code)
    PushMap pm;
    pm.code = pd.x < 0 ? -1 : pd.w;
    const int n = P.n, g = P.g, m = P.m;
    const int sgn = pd.z ? -1 : 1, k0 = pd.z ? n - 1 : 0;     // k -> k0 + sgn * k along the destination edge
    const unsigned tile = (unsigned)max(pd.x, 0) * m * m;
    if (pd.y < 2) {
        const int dj = pd.y == 0 ? g + n + d : g - 1 - d;
        pm.base = tile + (unsigned)dj * m + g + k0;
        pm.stride = sgn;
    } else {
        const int di = pd.y == 2 ? g + n + d : g - 1 - d;
        pm.base = tile + (unsigned)(g + k0) * m + di;
        pm.stride = sgn * m;
    }
    return pm;
}

template <typename R>
__device__ __forceinline__ void march_push_store(const StageParams<R>& P, const PushMap& pm, int k, R h, R u, R v) {
    const unsigned o = pm.base + (unsigned)(k * pm.stride);
    const R x = (pm.code & 1) ? v : u, y = (pm.code & 1) ? u : v;
    P.ho[o] = h;
    P.uo[o] = (pm.code & 2) ? -x : x;
    P.vo[o] = (pm.code & 4) ? -y : y;
}

// x direction, tendencies and RK combination of one row segment (shared by the LDG-fed and the
// TMA-fed kernels).  Each lane computes the half slopes of its own V cells and the V faces on
// their left; the neighbour's slope and the right-most face come by shuffle, so no slope or flux
// is computed twice within a warp (and both sides of a face see one value).
template <typename R, int V, int LIM, bool UNIT = false>
__device__ __forceinline__ void march_row_update(const StageParams<R>& P, const MarchConst<R, V>& K,
                                                 const RowDerived<R, V>& Dcur, const RowDerived<R, V>& Dnext,
                                                 const R (&hc)[V], const R (&uc)[V], const R (&vc)[V], const R (&dG)[V],
                                                 const R (&c1m)[V], const R (&Bnm)[V], R Yc, R ayc, bool read_q0,
                                                 const R (&q0h)[V], const R (&q0u)[V], const R (&q0v)[V],
                                                 R (&oh)[V], R (&ou)[V], R (&ov)[V]) {
    const R hl = __shfl_up_sync(0xffffffffu, hc[V - 1], 1);
    const R hr = __shfl_down_sync(0xffffffffu, hc[0], 1);
    R sx[V], F[V + 1], c2e[V + 2], Bne[V + 2];
#pragma unroll
    for (int c = 0; c < V; ++c) {
        const R hm_ = c == 0 ? hl : hc[c > 0 ? c - 1 : 0];
        const R hp_ = c == V - 1 ? hr : hc[c < V - 1 ? c + 1 : V - 1];
        sx[c] = march_half_slope<R, LIM>(hc[c] - hm_, hp_ - hc[c]);
    }
    const R sxl = __shfl_up_sync(0xffffffffu, sx[V - 1], 1);
    const R m1l = __shfl_up_sync(0xffffffffu, Dcur.m1[V - 1], 1);
    F[0] = march_flux4<R>(m1l, Dcur.m1[0], hl, sxl, hc[0], sx[0]);
#pragma unroll
    for (int c = 1; c < V; ++c) F[c] = march_flux4<R>(Dcur.m1[c - 1], Dcur.m1[c], hc[c - 1], sx[c - 1], hc[c], sx[c]);
    F[V] = __shfl_down_sync(0xffffffffu, F[0], 1);
    c2e[0] = __shfl_up_sync(0xffffffffu, Dcur.c2[V - 1], 1);
    c2e[V + 1] = __shfl_down_sync(0xffffffffu, Dcur.c2[0], 1);
    Bne[0] = __shfl_up_sync(0xffffffffu, Dcur.Bn[V - 1], 1);
    Bne[V + 1] = __shfl_down_sync(0xffffffffu, Dcur.Bn[0], 1);
#pragma unroll
    for (int c = 0; c < V; ++c) {
        c2e[c + 1] = Dcur.c2[c];
        Bne[c + 1] = Dcur.Bn[c];
    }

// tendencies and the RK combination
    const R ay = ayc, Yay = Yc * ayc;
#pragma unroll
    for (int c = 0; c < V; ++c) {
        // every factor b, dt, 1/delta, 1/a^2 is already inside isg, q and fsg (MarchConst)
        // UNIT: the first stage of every scheme, (a, b) = (0, 1) -- no q0, no multiplication by b
        oh[c] = (UNIT ? hc[c] : P.b * hc[c]) - ((F[c + 1] - F[c]) + dG[c]) * Dcur.isg[c];
        ou[c] = UNIT ? uc[c] : P.b * uc[c];
        ov[c] = UNIT ? vc[c] : P.b * vc[c];
        {   // advection only: kq == 0 switches the tendencies off; a branch here would cut the row body into
            // one basic block per column and keep the scheduler from interleaving the V dependency chains
            const R eta = ((c2e[c + 2] - c2e[c]) - (Dnext.c1[c] - c1m[c])) + Dcur.fsg[c];
            const R T1 = eta * vc[c] - (Bne[c + 2] - Bne[c]);
            const R T2 = -eta * uc[c] - (Dnext.Bn[c] - Bnm[c]);
            const R w12 = K.Xax[c] * Yay;
            ou[c] += Dcur.q[c] * (K.ax[c] * T1 + w12 * T2);
            ov[c] += Dcur.q[c] * (w12 * T1 + ay * T2);
        }
        if (!UNIT && read_q0) {
            oh[c] += P.a * q0h[c];
            ou[c] += P.a * q0u[c];
            ov[c] += P.a * q0v[c];
        }
    }
}

// SUMS: 0 = none; 1 = the launch also leaves mass / energy of its OUTPUT cells in P.diag_out (K3 fused into K2, new-state
// semantics; costs registers and ~16 instructions per cell); 2 = of its INPUT cells (taken in march_derive, ~6
// instructions per cell and no registers: the cheap way to monitor conservation, one stage behind).
//
// march_item is the whole stage for ONE work item (tile, strip, chunk), run by one warp; the two kernels below call it:
// stage_march_kernel (one stage per launch, one item per warp) and step_march_kernel (a whole time step per launch).
// PDL: the caller was launched as a programmatic dependent and the state may be touched only after the grid dependency.
template <typename R, int V, int LIM, bool PUSH, bool UNIT, int SUMS, bool PDL>
__device__ __forceinline__ void march_item(const StageParams<R>& P, const MarchGeom& G, const int item, const int lane,
                                           double& dmass, double& denergy2) {
    const int chunk = item % G.nchunks;
    const int strip = (item / G.nchunks) % G.nstrips;
    const int t = item / (G.nchunks * G.nstrips);
    const int j0 = G.ylo + strip * G.rows;
    const int j1 = min(j0 + G.rows, G.yhi);
    const int m = P.m, g = P.g;
    const TileCoord tc = cs_tile_coord(cs_tile_id(P, t), P.p);
    const int grp = G.grp0 + chunk * CS_MARCH_OUT_LANES + lane - 1;
    const bool lane_out = lane >= 1 && lane <= CS_MARCH_OUT_LANES && grp < G.ngroups;
    const int col0 = min(max(grp, 0), G.ngroups - 1) * V;
    const R* yt = P.xt + tc.ty * P.n;
    const R* ayt = P.ainv + tc.ty * P.n;
    const R* cf = P.cf[tc.face];
    const bool has_topo = P.topo != nullptr, read_q0 = !UNIT && P.a != R(0);

    MarchConst<R, V> K;
    bool om[V];
    bool all_out = true;
    {
        const R kf = P.two_omega / P.r2d;
        K.kq = P.ia2 * P.b * P.dt * P.r2d;
        if (P.advect_only) K.kq = R(0);
        K.kisg = R(0.5) * P.ia2 * P.b * P.dt * P.rd;    // 1/2: fluxes are carried times two
        K.kf2 = kf * cf[2];
#pragma unroll
        for (int c = 0; c < V; ++c) {
            K.X[c] = P.xt[tc.tx * P.n + col0 + c];
            K.ax[c] = P.ainv[tc.tx * P.n + col0 + c];
            K.a2A[c] = P.a2 * (R(1) + K.X[c] * K.X[c]);
            K.cfx[c] = kf * (cf[0] + cf[1] * K.X[c]);
            K.Xax[c] = K.X[c] * K.ax[c];
            const int i = col0 + c - g;
            om[c] = lane_out && i >= G.xlo && i < G.xhi;
            all_out = all_out && om[c];
        }
    }
    // fused halo: only the first / last chunk of a row holds columns within g of the tile's W / E edge.
    // A lane's edge columns share one ghost mapping up to the depth: depth d adds d * ew_dstep elements.
    const bool edge_warp = PUSH && (chunk == 0 || chunk == G.nchunks - 1);
    PushMap ew;
    int ew_dstep = 0;
    ew.code = -1;
    if constexpr (PUSH) {
        const int i0 = col0 - g, i1 = col0 + V - 1 - g;              // interior index of the first / last column
        if (edge_warp && lane_out && (i0 < g || i1 >= P.n - g)) {
            const int e = i0 < g ? 3 : 2;
            ew = march_push_map<R>(P, t, e, 0);
            ew_dstep = (int)(march_push_map<R>(P, t, e, 1).base - ew.base);
        }
    }
    const unsigned tbase = (unsigned)t * m * m + col0;   // shards are < 2^31 elements (checked at create)
    using IO = VecIO<R, V>;

    // Launched as a programmatic dependent: everything above (indices, metric constants from tables no kernel writes)
    // may run while the previous kernel in the stream drains; the state is touched only after this point.  The
    // trigger lets the next kernel's blocks be scheduled as this grid's blocks retire.
    if constexpr (PDL) {
        cudaGridDependencySynchronize();
        cudaTriggerProgrammaticLaunchCompletion();
    }

    // ---- prologue: rows J0-2 .. J0+1 --------------------------------------------------------
    const int J0 = j0 + g, J1 = j1 + g;
    R hc[V], hp[V], nh[V], uc[V], vc[V], nu[V], nv[V], nt[V];
    R c1m[V], Bnm[V], Gs[V], syc[V];
    RowDerived<R, V> Dc;
    R Yc = yt[J0], ayc = ayt[J0];                         // tan(beta) and 1/(1+Y^2) of the current row
    R Yn = yt[J0 + 1], ayn = ayt[J0 + 1];                 // ... and of the next one (loaded a row ahead)
    const int pf_rows = G.prefetch;                       // L2 prefetch distance in rows (0 = off)
    const bool pf_lane = (lane * (int)(V * sizeof(R))) % 128 == 0;   // one lane per 128-byte line
    {
        R hmm[V], hm1[V], um[V], vm[V], tm[V], tcur[V];
        IO::ld(P.hq + tbase + (unsigned)(J0 - 2) * m, hmm);
        IO::ld(P.hq + tbase + (unsigned)(J0 - 1) * m, hm1);
        IO::ld(P.hq + tbase + (unsigned)J0 * m, hc);
        IO::ld(P.hq + tbase + (unsigned)(J0 + 1) * m, hp);
        IO::ld(P.uq + tbase + (unsigned)(J0 - 1) * m, um);
        IO::ld(P.vq + tbase + (unsigned)(J0 - 1) * m, vm);
        IO::ld(P.uq + tbase + (unsigned)J0 * m, uc);
        IO::ld(P.vq + tbase + (unsigned)J0 * m, vc);
#pragma unroll
        for (int c = 0; c < V; ++c) tm[c] = tcur[c] = nt[c] = R(0);
        if (has_topo) {
            IO::ld(P.topo + tbase + (unsigned)(J0 - 1) * m, tm);
            IO::ld(P.topo + tbase + (unsigned)J0 * m, tcur);
            IO::ld(P.topo + tbase + (unsigned)(J0 + 1) * m, nt);
        }
        IO::ld(P.hq + tbase + (unsigned)min(J0 + 2, m - 1) * m, nh);
        IO::ld(P.uq + tbase + (unsigned)(J0 + 1) * m, nu);
        IO::ld(P.vq + tbase + (unsigned)(J0 + 1) * m, nv);
        RowDerived<R, V> Dm;
        march_derive<R, V>(P, K, yt[J0 - 1], ayt[J0 - 1], hm1, um, vm, tm, Dm);
        if constexpr (SUMS == 2) march_derive<R, V>(P, K, Yc, ayc, hc, uc, vc, tcur, Dc, InSums<R, V>{om, true, dmass, denergy2});
        else march_derive<R, V>(P, K, Yc, ayc, hc, uc, vc, tcur, Dc);
#pragma unroll
        for (int c = 0; c < V; ++c) {
            const R sym = march_half_slope<R, LIM>(hm1[c] - hmm[c], hc[c] - hm1[c]);
            syc[c] = march_half_slope<R, LIM>(hc[c] - hm1[c], hp[c] - hc[c]);
            Gs[c] = march_flux4<R>(Dm.m2[c], Dc.m2[c], hm1[c], sym, hc[c], syc[c]);
            c1m[c] = Dm.c1[c];
            Bnm[c] = Dm.Bn[c];
        }
    }

    // ---- march ---------------------------------------------------------------------------------
    // One row per call.  `Dcur` holds the derived quantities of row J, `Dnext` receives those of
    // row J+1; the loop calls the body twice with the two sets swapped, so they never get copied.
    auto row = [&](RowDerived<R, V>& Dcur, RowDerived<R, V>& Dnext, const int J) {
        // loads for the next call (rows J+3 / J+2) and the point-wise q0 of this row
        R ph[V], pu[V], pv[V], pt[V], q0h[V], q0u[V], q0v[V];
        const unsigned oJ = tbase + (unsigned)J * m;
        const unsigned o2 = tbase + (unsigned)min(J + 2, m - 1) * m;
        IO::ld(P.hq + tbase + (unsigned)min(J + 3, m - 1) * m, ph);
        IO::ld(P.uq + o2, pu);
        IO::ld(P.vq + o2, pv);
        if (has_topo) IO::ld(P.topo + o2, pt);
        if (read_q0) {
            IO::ld(P.h0 + oJ, q0h);
            IO::ld(P.u0 + oJ, q0u);
            IO::ld(P.v0 + oJ, q0v);
        }
        // Pull the lines that will be loaded pf_rows rows from now into L2 (no registers involved): the
        // register prefetch above is one row deep, which covers an L2 hit but not a trip to HBM.
        if (pf_rows > 0 && pf_lane) {
            const unsigned op = tbase + (unsigned)min(J + 3 + pf_rows, m - 1) * m;
            march_prefetch_l2(P.hq + op);
            march_prefetch_l2(P.uq + op);
            march_prefetch_l2(P.vq + op);
            if (has_topo) march_prefetch_l2(P.topo + op);
            if (read_q0) {
                const unsigned oq = tbase + (unsigned)min(J + pf_rows, m - 1) * m;
                march_prefetch_l2(P.h0 + oq);
                march_prefetch_l2(P.u0 + oq);
                march_prefetch_l2(P.v0 + oq);
            }
        }

        // row J+1: derived quantities, y-slope and the north-face flux of row J
        const R Yn2 = yt[min(J + 2, m - 1)], ayn2 = ayt[min(J + 2, m - 1)];
        if constexpr (SUMS == 2) march_derive<R, V>(P, K, Yn, ayn, hp, nu, nv, nt, Dnext, InSums<R, V>{om, J + 1 < J1, dmass, denergy2});
        else march_derive<R, V>(P, K, Yn, ayn, hp, nu, nv, nt, Dnext);
        R dG[V];
#pragma unroll
        for (int c = 0; c < V; ++c) {
            const R syp = march_half_slope<R, LIM>(hp[c] - hc[c], nh[c] - hp[c]);
            const R Gn = march_flux4<R>(Dcur.m2[c], Dnext.m2[c], hc[c], syc[c], hp[c], syp);
            dG[c] = Gn - Gs[c];
            Gs[c] = Gn;
            syc[c] = syp;
        }

        R oh[V], ou[V], ov[V];
        march_row_update<R, V, LIM, UNIT>(P, K, Dcur, Dnext, hc, uc, vc, dG, c1m, Bnm, Yc, ayc, read_q0, q0h, q0u, q0v, oh, ou, ov);
        // (per-lane branch on purpose: a warp-uniform choice of the store path -- vector stores predicated on the output
        // lanes, the masked path only in the edge chunks -- removes the divergence of the two halo lanes and 8 issue
        // slots per row, and ran stage 1 at 89 us instead of 82 and stages 2 / 3 at 110 instead of 89: DESIGN.md 4.2)
        if (all_out) {
            IO::st(P.ho + oJ, oh);
            IO::st(P.uo + oJ, ou);
            IO::st(P.vo + oJ, ov);
        } else {
#pragma unroll
            for (int c = 0; c < V; ++c)
                if (om[c]) P.ho[oJ + c] = oh[c], P.uo[oJ + c] = ou[c], P.vo[oJ + c] = ov[c];
        }
        if constexpr (SUMS == 1) {
            R hsJ[V];
#pragma unroll
            for (int c = 0; c < V; ++c) hsJ[c] = R(0);
            if (has_topo) IO::ld(P.topo + oJ, hsJ);       // row J again: it was loaded two rows ago and sits in L1 / L2
            march_row_diag<R, V>(P, K, Dcur, Yc, oh, ou, ov, hsJ, om, dmass, denergy2);
        }
        if constexpr (PUSH) {
            const int j = J - g;
            if (j < g || j >= P.n - g) {                  // whole row lies in the S / N boundary band
                const PushMap pm = march_push_map<R>(P, t, j < g ? 1 : 0, j < g ? j : P.n - 1 - j);
                if (pm.code >= 0) {
#pragma unroll
                    for (int c = 0; c < V; ++c)
                        if (om[c]) march_push_store<R>(P, pm, col0 + c - g, oh[c], ou[c], ov[c]);
                }
            }
            if (edge_warp && ew.code >= 0) {
#pragma unroll
                for (int c = 0; c < V; ++c) {
                    const int i = col0 + c - g, d = i < g ? i : P.n - 1 - i;
                    if (om[c] && d < g) {
                        PushMap pm = ew;
                        pm.base += (unsigned)(d * ew_dstep);
                        march_push_store<R>(P, pm, j, oh[c], ou[c], ov[c]);
                    }
                }
            }
        }

        // slide the raw-row window one row north
#pragma unroll
        for (int c = 0; c < V; ++c) {
            c1m[c] = Dcur.c1[c];
            Bnm[c] = Dcur.Bn[c];
            hc[c] = hp[c], hp[c] = nh[c], nh[c] = ph[c];
            uc[c] = nu[c], vc[c] = nv[c], nu[c] = pu[c], nv[c] = pv[c];
            if (has_topo) nt[c] = pt[c];
        }
        Yc = Yn, ayc = ayn, Yn = Yn2, ayn = ayn2;
    };

    RowDerived<R, V> Dn;
    int J = J0;
#pragma unroll 1
    for (; J + 1 < J1; J += 2) {
        row(Dc, Dn, J);
        row(Dn, Dc, J + 1);
    }
    if (J < J1) row(Dc, Dn, J);
}
// One (mass, energy) partial per slot (a work item, or a warp of the step kernel); the warp that takes the last ticket
// adds the partials in index order, so the result does not depend on which warp finishes when (no floating-point
// atomics).  `scale_e`: SUMS == 1 carries twice the energy density, SUMS == 2 the density itself.
template <typename R>
__device__ __forceinline__ void march_finish_sums(const StageParams<R>& P, int slot, int nslots, int lane, double dmass,
                                                  double denergy2, double scale_e) {
    double pm = dmass * (double)P.darea, pe = denergy2 * (scale_e * (double)P.darea);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        pm += __shfl_down_sync(0xffffffffu, pm, o);
        pe += __shfl_down_sync(0xffffffffu, pe, o);
    }
    unsigned int ticket = 0;
    if (lane == 0) {
        P.partials[2 * slot] = pm;
        P.partials[2 * slot + 1] = pe;
        __threadfence();
        ticket = atomicAdd(P.ticket, 1u);
    }
    ticket = __shfl_sync(0xffffffffu, ticket, 0);
    if (ticket == (unsigned int)nslots - 1u) {
        __threadfence();
        double a = 0.0, b = 0.0;
        for (int k = lane; k < nslots; k += 32) {
            a += __ldcg(P.partials + 2 * k);              // other warps wrote these: read through L2, never a stale L1 line
            b += __ldcg(P.partials + 2 * k + 1);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            a += __shfl_down_sync(0xffffffffu, a, o);
            b += __shfl_down_sync(0xffffffffu, b, o);
        }
        if (lane == 0) {
            P.diag_out[0] = a, P.diag_out[1] = b;
            *P.ticket = 0;                                // ready for the next launch on this stream
        }
    }
}

// One stage per launch: one work item per warp, items flattened chunk-fastest into a 1-D grid, so consecutive warps
// take neighbouring column chunks of the same strip and no warp slot idles when nchunks % 4 != 0.
template <typename R, int V, int LIM, int MINB, bool PUSH, bool UNIT = false, int SUMS = 0>
__global__ void __launch_bounds__(CS_MARCH_THREADS, MINB)
stage_march_kernel(const __grid_constant__ StageParams<R> P, const __grid_constant__ MarchGeom G) {
    const int lane = threadIdx.x & 31;
    const int item = blockIdx.x * (CS_MARCH_THREADS / 32) + (threadIdx.x >> 5);
    if (item >= G.nitems) return;                         // whole warp leaves together
    double dmass = 0.0, denergy2 = 0.0;                   // SUMS: this thread's share of the conservation sums
    march_item<R, V, LIM, PUSH, UNIT, SUMS, true>(P, G, item, lane, dmass, denergy2);
    // (A push of the boundary cells into the neighbours' ghosts after the march -- march_push_item, as the step kernel
    // does -- was tried here as an option for cs_step's intermediate stages: slower when on, and its mere presence in this
    // kernel, switched off, cost the fp32 step 8 us.  DESIGN.md 4.2.)
    if constexpr (SUMS != 0) march_finish_sums<R>(P, item, G.nitems, lane, dmass, denergy2, SUMS == 1 ? 0.5 : 1.0);
}

// ---------------------------------------------------------------------------------------------------------------
// CS step kernel: one launch per time step.
//
// A step of the one-stage-per-launch path is 6 launches (3 stages, 3 ghost fills): at C1536 fp32 the gaps between
// them and the short fill kernels cost 30 of 282 microseconds, and on small spheres they are most of the step.  Here
// one cooperative grid runs all stages: a warp marches its work items, then copies the boundary cells it has just
// written into the ghost cells of the neighbouring tiles that mirror them (the same index map and signed permutation as
// the ghost-fill kernel, but only where it matters: edge strips, in a pass after the march so that the marching loop
// keeps its registers), and the whole grid meets at a barrier before the next stage.  After the last stage the
// table-driven fill runs once over every ghost cell, so the state the caller gets back has the same ghosts, corner
// blocks included, as after cs_halo_fill_state.  Needs every neighbour tile on this device (a whole-sphere context).
// ---------------------------------------------------------------------------------------------------------------
#include <cooperative_groups.h>

template <typename R>
struct StepParams {
    StageParams<R> st[3];
    int nstages;                     // 2 or 3
    const int32_t* halo_dst;         // the context's ghost table (all local)
    const int32_t* halo_src;
    const uint8_t* halo_code;
    long long halo_count;
    int debug;                       // CS_STEP_DEBUG (timing experiments only): 1 no push, 2 no final fill, 4 no barriers,
                                     // 8 first stage only, 16 record the trace below
    unsigned long long* trace;       // [gridDim.x][8] globaltimer at: start, end of march 1, 2, 3, after each barrier
};

__device__ __forceinline__ unsigned long long cs_globaltimer() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
#define CS_STEP_TRACE(slot)                                                                  \
    do {                                                                                     \
        if ((S.debug & 16) && threadIdx.x == 0) S.trace[blockIdx.x * 8 + (slot)] = cs_globaltimer(); \
    } while (0)

// Boundary cells of one finished work item -> the neighbours' edge ghosts.  All lanes take part: lanes read cells other
// lanes of the warp wrote (after __syncwarp; through L2, which is where the stores went).
template <typename R, int V>
__device__ __forceinline__ void march_push_item(const StageParams<R>& P, const MarchGeom& G, const int item, const int lane) {
    const int chunk = item % G.nchunks;
    const int strip = (item / G.nchunks) % G.nstrips;
    const int t = item / (G.nchunks * G.nstrips);
    const int n = P.n, g = P.g, m = P.m;
    const int j0 = G.ylo + strip * G.rows, j1 = min(j0 + G.rows, G.yhi);
    // interior columns this chunk's output lanes own
    const int i0 = max(G.xlo, (G.grp0 + chunk * CS_MARCH_OUT_LANES) * V - g);
    const int i1 = min(G.xhi, (G.grp0 + (chunk + 1) * CS_MARCH_OUT_LANES) * V - g);
    const unsigned tile = (unsigned)t * m * m;
    const bool south = j0 < g, north = j1 > n - g, west = i0 < g, east = i1 > n - g;
    if (!(south || north || west || east)) return;        // most items touch no tile edge
    __syncwarp();
    for (int d = 0; d < g; ++d) {
        if (south && d >= j0 && d < j1) {
            const PushMap pm = march_push_map<R>(P, t, 1, d);
            if (pm.code >= 0)
                for (int i = i0 + lane; i < i1; i += 32) {
                    const unsigned o = tile + (unsigned)(d + g) * m + i + g;
                    march_push_store<R>(P, pm, i, __ldcg(P.ho + o), __ldcg(P.uo + o), __ldcg(P.vo + o));
                }
        }
        const int jn = n - 1 - d;
        if (north && jn >= j0 && jn < j1) {
            const PushMap pm = march_push_map<R>(P, t, 0, d);
            if (pm.code >= 0)
                for (int i = i0 + lane; i < i1; i += 32) {
                    const unsigned o = tile + (unsigned)(jn + g) * m + i + g;
                    march_push_store<R>(P, pm, i, __ldcg(P.ho + o), __ldcg(P.uo + o), __ldcg(P.vo + o));
                }
        }
        if (west && d >= i0 && d < i1) {
            const PushMap pm = march_push_map<R>(P, t, 3, d);
            if (pm.code >= 0)
                for (int j = j0 + lane; j < j1; j += 32) {
                    const unsigned o = tile + (unsigned)(j + g) * m + d + g;
                    march_push_store<R>(P, pm, j, __ldcg(P.ho + o), __ldcg(P.uo + o), __ldcg(P.vo + o));
                }
        }
        const int ie = n - 1 - d;
        if (east && ie >= i0 && ie < i1) {
            const PushMap pm = march_push_map<R>(P, t, 2, d);
            if (pm.code >= 0)
                for (int j = j0 + lane; j < j1; j += 32) {
                    const unsigned o = tile + (unsigned)(j + g) * m + ie + g;
                    march_push_store<R>(P, pm, j, __ldcg(P.ho + o), __ldcg(P.uo + o), __ldcg(P.vo + o));
                }
        }
    }
}

// UNIT1: the first stage uses the (a, b) = (0, 1) build (fp64); SUMS1: 0, or 2 = the launch also leaves mass / energy
// of the state the step starts from in st[0].diag_out.
template <typename R, int V, int LIM, int MINB, bool UNIT1, int SUMS1>
__global__ void __launch_bounds__(CS_MARCH_THREADS, MINB)
step_march_kernel(const __grid_constant__ StepParams<R> S, const __grid_constant__ MarchGeom G) {
    cooperative_groups::grid_group grid = cooperative_groups::this_grid();
    const int lane = threadIdx.x & 31;
    const int warp = blockIdx.x * (CS_MARCH_THREADS / 32) + (threadIdx.x >> 5);
    const int nwarps = gridDim.x * (CS_MARCH_THREADS / 32);
    double dmass = 0.0, denergy = 0.0;
    CS_STEP_TRACE(0);
#pragma unroll 1
    for (int item = warp; item < G.nitems; item += nwarps) {
        march_item<R, V, LIM, false, UNIT1, SUMS1, false>(S.st[0], G, item, lane, dmass, denergy);
        if (!(S.debug & 1)) march_push_item<R, V>(S.st[0], G, item, lane);
    }
    if constexpr (SUMS1 == 2) march_finish_sums<R>(S.st[0], warp, nwarps, lane, dmass, denergy, 1.0);
    double unused0 = 0.0, unused1 = 0.0;
    __syncthreads();
    CS_STEP_TRACE(1);
    if (S.nstages > 1 && !(S.debug & 8)) {
        __threadfence();
        if (!(S.debug & 4)) grid.sync();
        CS_STEP_TRACE(2);
#pragma unroll 1
        for (int item = warp; item < G.nitems; item += nwarps) {
            march_item<R, V, LIM, false, false, 0, false>(S.st[1], G, item, lane, unused0, unused1);
            if (!(S.debug & 1)) march_push_item<R, V>(S.st[1], G, item, lane);
        }
    }
    __syncthreads();
    CS_STEP_TRACE(3);
    if (S.nstages > 2 && !(S.debug & 8)) {
        __threadfence();
        if (!(S.debug & 4)) grid.sync();
        CS_STEP_TRACE(4);
#pragma unroll 1
        for (int item = warp; item < G.nitems; item += nwarps) {
            march_item<R, V, LIM, false, false, 0, false>(S.st[2], G, item, lane, unused0, unused1);
            if (!(S.debug & 1)) march_push_item<R, V>(S.st[2], G, item, lane);
        }
    }
    // every ghost cell of the new state, corner blocks included, exactly as the ghost-fill kernel leaves them
    __syncthreads();
    CS_STEP_TRACE(5);
    if (S.debug & 2) return;
    __threadfence();
    if (!(S.debug & 4)) grid.sync();
    CS_STEP_TRACE(6);
    const StageParams<R>& L = S.st[S.nstages > 2 ? 2 : S.nstages - 1];
    R* h = L.ho;                                           // sources are interior cells, targets ghost cells of the same arrays
    R* u = L.uo;
    R* v = L.vo;
    for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < S.halo_count; k += (long long)gridDim.x * blockDim.x) {
        const int32_t d = S.halo_dst[k], s = S.halo_src[k];
        const int code = S.halo_code[k];
        const R a = __ldcg(u + s), b = __ldcg(v + s);
        h[d] = __ldcg(h + s);
        const R x = (code & 1) ? b : a, y = (code & 1) ? a : b;
        u[d] = (code & 2) ? -x : x;
        v[d] = (code & 4) ? -y : y;
    }
    __syncthreads();
    CS_STEP_TRACE(7);
}
