// CS_KERNEL_TMA -- the register-marching stage fed by a TMA shared-memory ring.
//
// Same arithmetic as stage_march_kernel (march_derive / march_row_update), but the raw rows
// (h, u, v of the stage input, the point-wise q0, the topography) never pass through registers on
// their way in: lane 0 of every warp issues bulk asynchronous copies (cp.async.bulk, the TMA
// engine; SASS UBLKCP) of whole row segments -- 32 lanes x 8 or 16 bytes per array -- into a ring
// of row slots in shared memory, each slot guarded by an mbarrier that counts the bytes landed.
// Lanes read their bytes from the slot when the row is due.  The ring keeps (ring - 3) rows per
// warp in flight independent of the register budget, which is what hides the HBM latency for a
// stencil that carries this much state per thread.
// Rings are warp-private: no block-level barrier; a warp re-arms a slot right after its last read
// of it (program order + __syncwarp is the only "empty" signal needed).
//
// Bulk copies need 16-byte aligned addresses and sizes.  A lane owns LB = V * sizeof(R) bytes (8 or
// 16), so a warp's segment may start 8 bytes off a 16-byte boundary: the copy then starts 8 bytes
// early, and data sit in shared memory at the same offset modulo 16 as in global memory.
#pragma once

#include "cs_stage_march.cuh"

#define CS_TMA_MAX_RING 12

namespace cs_tma {

__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
                 "l"(src), "r"(bytes), "r"(bar)
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" ::"r"(bar),
        "r"(parity)
        : "memory");
}

// What one warp copies per row: byte geometry of its segment and which rows carry which arrays.
struct RowPlan {
    const char* src[7];       // h, u, v, h0, u0, v0, topo at (tile, row 0, segment start): add row * pitch
    uint32_t pitch;           // bytes per row
    uint32_t bytes;           // bytes per copy (multiple of 16)
    uint32_t dst_off;         // byte offset of the copy inside an array region
    uint32_t astride;         // bytes per array region
    uint32_t slot_bytes;
    int J0, J1;               // u, v, topo for rows [J0-1, J1]; q0 for rows [J0, J1)
    int a_topo;               // region index of the topography (3 or 6); q0 regions are 3, 4, 5
    bool read_q0, has_topo;
};

// Arm the slot's barrier and launch the row's copies.  One lane calls this, once per row.
__device__ __noinline__ void issue_row(const RowPlan& pl, int r, uint32_t slot_addr, uint32_t bar) {
    const bool uv = r >= pl.J0 - 1 && r <= pl.J1, q0 = pl.read_q0 && r >= pl.J0 && r < pl.J1;
    const int count = 1 + (uv ? 2 : 0) + (uv && pl.has_topo ? 1 : 0) + (q0 ? 3 : 0);
    const size_t ro = (size_t)r * pl.pitch;
    const uint32_t dst = slot_addr + pl.dst_off;
    mbar_expect_tx(bar, pl.bytes * count);
    bulk_g2s(dst, pl.src[0] + ro, pl.bytes, bar);
    if (uv) {
        bulk_g2s(dst + 1 * pl.astride, pl.src[1] + ro, pl.bytes, bar);
        bulk_g2s(dst + 2 * pl.astride, pl.src[2] + ro, pl.bytes, bar);
        if (pl.has_topo) bulk_g2s(dst + pl.a_topo * pl.astride, pl.src[6] + ro, pl.bytes, bar);
    }
    if (q0) {
        bulk_g2s(dst + 3 * pl.astride, pl.src[3] + ro, pl.bytes, bar);
        bulk_g2s(dst + 4 * pl.astride, pl.src[4] + ro, pl.bytes, bar);
        bulk_g2s(dst + 5 * pl.astride, pl.src[5] + ro, pl.bytes, bar);
    }
}

}  // namespace cs_tma

template <typename R, int V>
__host__ __device__ constexpr uint32_t cs_tma_astride() { return 32u * V * sizeof(R) + 16u; }

template <typename R, int V, int LIM, int MINB>
__global__ void __launch_bounds__(CS_MARCH_THREADS, MINB)
stage_tma_kernel(const __grid_constant__ StageParams<R> P, const __grid_constant__ MarchGeom G, const int ring) {
    constexpr uint32_t LB = V * sizeof(R);                // bytes per lane per array
    static_assert(LB == 8 || LB == 16, "a lane owns 8 or 16 bytes of every row");
    constexpr uint32_t AS = cs_tma_astride<R, V>();
    extern __shared__ __align__(128) unsigned char cs_tma_smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int item = blockIdx.x * (CS_MARCH_THREADS / 32) + warp;      // (tile, strip, chunk), chunk fastest
    if (item >= G.nitems) return;                         // whole warp leaves together
    const int chunk = item % G.nchunks;
    const int strip = (item / G.nchunks) % G.nstrips;
    const int t = item / (G.nchunks * G.nstrips);
    const int j0 = G.ylo + strip * G.rows;
    const int j1 = min(j0 + G.rows, G.yhi);
    const int m = P.m, g = P.g;
    const TileCoord tc = cs_tile_coord(cs_tile_id(P, t), P.p);
    const int gs = G.grp0 + chunk * CS_MARCH_OUT_LANES - 1;             // column group of lane 0 (may be -1)
    const int grp = gs + lane;
    const bool lane_out = lane >= 1 && lane <= CS_MARCH_OUT_LANES && grp < G.ngroups;
    const int col0 = min(max(grp, 0), G.ngroups - 1) * V;
    const R* yt = P.xt + tc.ty * P.n;
    const R* ayt = P.ainv + tc.ty * P.n;
    const R* cf = P.cf[tc.face];
    const bool has_topo = P.topo != nullptr, read_q0 = P.a != R(0);

    MarchConst<R, V> K;
    bool om[V];
    bool all_out = true;
    {
        const R kf = P.two_omega / P.r2d;
        K.kq = P.advect_only ? R(0) : P.ia2 * P.b * P.dt * P.r2d;   // 0 switches the momentum tendencies off
        K.kisg = R(0.5) * P.ia2 * P.b * P.dt * P.rd;
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
    const unsigned obase = (unsigned)t * m * m + col0;                  // element offset of this lane's stores
    using IO = VecIO<R, V>;

    // ---- ring geometry ---------------------------------------------------------------------------
    const int narr = 3 + (read_q0 ? 3 : 0) + (has_topo ? 1 : 0);
    const int J0 = j0 + g, J1 = j1 + g;
    const int Jfirst = J0 - 2, Jlast = J1 + 1;                          // rows this strip touches
    cs_tma::RowPlan pl;
    pl.astride = AS;
    pl.slot_bytes = narr * AS;
    pl.J0 = J0, pl.J1 = J1, pl.read_q0 = read_q0, pl.has_topo = has_topo;
    pl.a_topo = read_q0 ? 6 : 3;
    pl.pitch = (uint32_t)m * sizeof(R);
    // lane l's bytes sit at region + lane_off + l * LB, lane_off = (gs * LB) mod 16, so that shared and
    // global addresses agree modulo 16
    const uint32_t lane_off = (uint32_t)(gs * (int)LB) & 15u;
    {
        const int gsc = max(gs, 0), gend = min(gs + 32, G.ngroups);
        const uint32_t b0 = (uint32_t)gsc * LB, b0a = b0 & ~15u;        // copy start in the row, aligned down
        const uint32_t b1 = ((uint32_t)gend * LB + 15u) & ~15u;         // copy end, aligned up (row pitch is 16-aligned)
        pl.bytes = b1 - b0a;
        pl.dst_off = lane_off + (uint32_t)(gsc - gs) * LB - (b0 - b0a);
        const size_t tile_off = (size_t)t * m * m * sizeof(R) + b0a;
        const void* bases[7] = {P.hq, P.uq, P.vq, P.h0, P.u0, P.v0, P.topo};
#pragma unroll
        for (int k = 0; k < 7; ++k) pl.src[k] = reinterpret_cast<const char*>(bases[k]) + tile_off;
    }
    const uint32_t slot_bytes = narr * AS;                              // hot-path copies (pl lives in local memory)
    const int a_topo = read_q0 ? 6 : 3;
    const uint32_t warp_bytes = ring * slot_bytes + CS_TMA_MAX_RING * 8;
    unsigned char* wbase = cs_tma_smem + (size_t)warp * warp_bytes;
    const uint32_t sbase = cs_tma::smem_addr(wbase);
    const uint32_t bbase = sbase + ring * slot_bytes;
    const unsigned char* lbase = wbase + lane_off + lane * LB;          // this lane's bytes in slot 0, region 0

    auto lds = [&](int slot, int arr, R (&o)[V]) {
        IO::ld(reinterpret_cast<const R*>(lbase + slot * slot_bytes + arr * AS), o);
    };

    if (lane == 0) {
        for (int s = 0; s < ring; ++s) cs_tma::mbar_init(bbase + s * 8, 1);
        cs_tma::fence_mbar_init();
    }
    __syncwarp();
    if (lane == 0)
        for (int q = 0; q < ring && Jfirst + q <= Jlast; ++q)
            cs_tma::issue_row(pl, Jfirst + q, sbase + q * slot_bytes, bbase + q * 8);

    // ---- prologue: rows J0-2 .. J0+1 (slots 0..3 of the first round) ----------------------------
    R hc[V], hp[V], uc[V], vc[V], nt[V];
    R c1m[V], Bnm[V], Gs[V], syc[V];
    RowDerived<R, V> Dc;
    R Yc = yt[J0], ayc = ayt[J0], Yn = yt[J0 + 1], ayn = ayt[J0 + 1];
    {
        R hmm[V], hm1[V], um[V], vm[V], tm[V], tcur[V];
#pragma unroll
        for (int c = 0; c < V; ++c) tm[c] = tcur[c] = nt[c] = R(0);
        cs_tma::mbar_wait(bbase + 0 * 8, 0);
        lds(0, 0, hmm);
        cs_tma::mbar_wait(bbase + 1 * 8, 0);
        lds(1, 0, hm1), lds(1, 1, um), lds(1, 2, vm);
        if (has_topo) lds(1, a_topo, tm);
        cs_tma::mbar_wait(bbase + 2 * 8, 0);
        lds(2, 0, hc), lds(2, 1, uc), lds(2, 2, vc);
        if (has_topo) lds(2, a_topo, tcur);
        cs_tma::mbar_wait(bbase + 3 * 8, 0);
        lds(3, 0, hp);
        RowDerived<R, V> Dm;
        march_derive<R, V>(P, K, yt[J0 - 1], ayt[J0 - 1], hm1, um, vm, tm, Dm);
        march_derive<R, V>(P, K, Yc, ayc, hc, uc, vc, tcur, Dc);
#pragma unroll
        for (int c = 0; c < V; ++c) {
            const R sym = march_half_slope<R, LIM>(hm1[c] - hmm[c], hc[c] - hm1[c]);
            syc[c] = march_half_slope<R, LIM>(hc[c] - hm1[c], hp[c] - hc[c]);
            Gs[c] = march_flux4<R>(Dm.m2[c], Dc.m2[c], hm1[c], sym, hc[c], syc[c]);
            c1m[c] = Dm.c1[c];
            Bnm[c] = Dm.Bn[c];
        }
        // rows J0-2 and J0-1 are dead: their slots take rows Jfirst + ring, Jfirst + ring + 1
        __syncwarp();
        if (lane == 0)
            for (int q = 0; q < 2; ++q)
                if (Jfirst + ring + q <= Jlast)
                    cs_tma::issue_row(pl, Jfirst + ring + q, sbase + q * slot_bytes, bbase + q * 8);
    }

    // slots of rows J, J+1, J+2 and the parity of row J+2's fill (ring >= 4)
    int s0 = 2, s1 = 3, s2 = 4 % ring;
    uint32_t p2 = (4 / ring) & 1;

    auto row = [&](RowDerived<R, V>& Dcur, RowDerived<R, V>& Dnext, const int J) {
        const R Yn2 = yt[J + 2], ayn2 = ayt[J + 2];       // for the next call; tables have m entries, J + 2 <= m - 1
        // row J+2 landed => so have rows J+1 and J (they were waited for in earlier calls)
        cs_tma::mbar_wait(bbase + s2 * 8, p2);
        R nh[V], nu[V], nv[V], q0h[V], q0u[V], q0v[V];
        lds(s2, 0, nh);
        lds(s1, 1, nu), lds(s1, 2, nv);
        if (has_topo) lds(s1, a_topo, nt);
        if (read_q0) lds(s0, 3, q0h), lds(s0, 4, q0u), lds(s0, 5, q0v);

        march_derive<R, V>(P, K, Yn, ayn, hp, nu, nv, nt, Dnext);
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
        march_row_update<R, V, LIM>(P, K, Dcur, Dnext, hc, uc, vc, dG, c1m, Bnm, Yc, ayc, read_q0, q0h, q0u, q0v, oh, ou, ov);
        const unsigned oJ = obase + (unsigned)J * m;
        if (all_out) {
            IO::st(P.ho + oJ, oh);
            IO::st(P.uo + oJ, ou);
            IO::st(P.vo + oJ, ov);
        } else {
#pragma unroll
            for (int c = 0; c < V; ++c)
                if (om[c]) P.ho[oJ + c] = oh[c], P.uo[oJ + c] = ou[c], P.vo[oJ + c] = ov[c];
        }
        // row J is dead: its slot takes row J + ring
        __syncwarp();
        if (lane == 0 && J + ring <= Jlast) cs_tma::issue_row(pl, J + ring, sbase + s0 * slot_bytes, bbase + s0 * 8);
#pragma unroll
        for (int c = 0; c < V; ++c) {
            c1m[c] = Dcur.c1[c];
            Bnm[c] = Dcur.Bn[c];
            hc[c] = hp[c], hp[c] = nh[c];
            uc[c] = nu[c], vc[c] = nv[c];
        }
        Yc = Yn, ayc = ayn, Yn = Yn2, ayn = ayn2;
        s0 = s1, s1 = s2;
        if (++s2 == ring) s2 = 0, p2 ^= 1u;
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
