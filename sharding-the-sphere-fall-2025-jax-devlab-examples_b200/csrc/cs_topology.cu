// Cube topology on the host: the 12 seams, the ghost-cell resolver, the gather
// tables of a context and the send/receive plan of a shard.
//
// Conventions come from the reference (face numbering, edge names, pairings and
// R/T flags: JAX-DevLab-Examples.py:114-139; ghost <- neighbour interior:
// JAX-DevLab-Examples.py:183-195).  The frames below are the unique right-handed
// assignment of the six faces to cube normals that reproduces every flag.
#include <algorithm>
#include <cstdio>

#include "cs_internal.cuh"

namespace {

// (normal, local x, local y)
const int FR[6][3][3] = {
    {{0, 0, 1}, {1, 0, 0}, {0, 1, 0}},   {{0, 1, 0}, {-1, 0, 0}, {0, 0, 1}},
    {{-1, 0, 0}, {0, -1, 0}, {0, 0, 1}}, {{0, -1, 0}, {1, 0, 0}, {0, 0, 1}},
    {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}},   {{0, 0, -1}, {-1, 0, 0}, {0, 1, 0}},
};

inline int dot(const int* a, const int* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }

// outward in-plane direction o and along-edge direction t of (face, edge)
void edge_vectors(int f, int e, int o[3], int t[3]) {
    const int* x = FR[f][1];
    const int* y = FR[f][2];
    for (int k = 0; k < 3; ++k) {
        o[k] = e == 0 ? y[k] : e == 1 ? -y[k] : e == 2 ? x[k] : -x[k];
        t[k] = e < 2 ? x[k] : y[k];
    }
}

}  // namespace

SeamInfo cs_seam(int f, int e) {
    int oa[3], ta[3];
    edge_vectors(f, e, oa, ta);
    SeamInfo s{-1, -1, 0, 0};
    for (int k = 0; k < 6; ++k)
        if (FR[k][0][0] == oa[0] && FR[k][0][1] == oa[1] && FR[k][0][2] == oa[2]) s.fb = k;
    int ob[3], tb[3];
    for (int k = 0; k < 4; ++k) {
        edge_vectors(s.fb, k, ob, tb);
        if (ob[0] == FR[f][0][0] && ob[1] == FR[f][0][1] && ob[2] == FR[f][0][2]) {
            s.eb = k;
            break;
        }
    }
    edge_vectors(s.fb, s.eb, ob, tb);
    s.rev = dot(ta, tb) < 0;
    // unfold the neighbour's axes into this face's plane: S[:, col] = image of (x_b, y_b)[col]
    int S[2][2];
    for (int col = 0; col < 2; ++col) {
        const int* w = FR[s.fb][1 + col];
        int img[3];
        for (int k = 0; k < 3; ++k) img[k] = dot(w, ta) * ta[k] - dot(w, ob) * oa[k];
        S[0][col] = dot(img, FR[f][1]);
        S[1][col] = dot(img, FR[f][2]);
    }
    int swap = S[0][0] == 0;
    int s1 = swap ? S[0][1] : S[0][0];
    int s2 = swap ? S[1][0] : S[1][1];
    s.code = swap | ((s1 < 0) << 1) | ((s2 < 0) << 2);
    return s;
}

void cs_resolve(int face, int64_t J, int64_t I, int N, int* fb, int64_t* Jb, int64_t* Ib, int* code) {
    bool xo = I < 0 || I >= N, yo = J < 0 || J >= N;
    if (xo && yo) {   // cube vertex block: constant extrapolation of the E/W ghost column in y
        J = std::min<int64_t>(std::max<int64_t>(J, 0), N - 1);
        yo = false;
    }
    if (!xo && !yo) {
        *fb = face, *Jb = J, *Ib = I, *code = 0;
        return;
    }
    int e;
    int64_t d, k;
    if (xo) {
        e = I >= N ? 2 : 3;
        d = I >= N ? I - N : -1 - I;
        k = J;
    } else {
        e = J >= N ? 0 : 1;
        d = J >= N ? J - N : -1 - J;
        k = I;
    }
    SeamInfo s = cs_seam(face, e);
    if (s.rev) k = N - 1 - k;
    *fb = s.fb;
    *code = s.code;
    switch (s.eb) {
        case 0: *Jb = N - 1 - d, *Ib = k; break;
        case 1: *Jb = d, *Ib = k; break;
        case 2: *Jb = k, *Ib = N - 1 - d; break;
        default: *Jb = k, *Ib = d; break;
    }
}

namespace {

// enumerate the ghost cells of global tile t in a fixed order
template <typename F>
void for_each_ghost(int t, int N, int p, int g, F&& fn) {
    const int n = N / p, m = n + 2 * g;
    const int face = t / (p * p), ty = (t % (p * p)) / p, tx = t % p;
    SeamInfo seams[4];
    for (int e = 0; e < 4; ++e) seams[e] = cs_seam(face, e);
    for (int jt = 0; jt < m; ++jt)
        for (int it = 0; it < m; ++it) {
            if (jt >= g && jt < g + n && it >= g && it < g + n) {
                it = g + n - 1;   // skip the interior of this row
                continue;
            }
            int fb, code;
            int64_t Jb, Ib;
            cs_resolve(face, (int64_t)ty * n + jt - g, (int64_t)tx * n + it - g, N, &fb, &Jb, &Ib, &code);
            int ts = fb * p * p + (int)(Jb / n) * p + (int)(Ib / n);
            int32_t src_local = (int32_t)((Jb % n + g) * m + (Ib % n + g));
            fn(jt, it, ts, src_local, code);
        }
}

template <typename T>
int upload(cs_ctx* ctx, T** dptr, const std::vector<T>& v) {
    if (*dptr) cudaFree(*dptr), *dptr = nullptr;
    if (v.empty()) return CS_OK;
    if (cudaMalloc((void**)dptr, v.size() * sizeof(T)) != cudaSuccess)
        return cs_fail(ctx, CS_E_NOMEM, "cudaMalloc of %zu bytes failed", v.size() * sizeof(T));
    CS_CUDA(ctx, cudaMemcpy(*dptr, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice));
    return CS_OK;
}

}  // namespace

int cs_build_halo(cs_ctx* ctx) {
    const int m = ctx->m;
    const int64_t tile_elems = (int64_t)m * m;
    ctx->halo.clear();
    for (int k = 0; k < ctx->T; ++k) {
        const int t = ctx->order[ctx->t0 + k];
        const int64_t base = (int64_t)k * tile_elems;
        for_each_ghost(t, ctx->N, ctx->p, ctx->g, [&](int jt, int it, int ts, int32_t src_local, int code) {
            HaloEntry e;
            e.dst = base + (int64_t)jt * m + it;
            e.src_tile = ts;
            e.src_local = src_local;
            e.code = code;
            const int li = ctx->local_index(ts);
            e.src = li >= 0 ? (int64_t)li * tile_elems + src_local : -1;
            ctx->halo.push_back(e);
        });
    }
    std::vector<int32_t> dst, src;
    std::vector<uint8_t> code;
    for (const HaloEntry& e : ctx->halo)
        if (e.src >= 0) {
            dst.push_back((int32_t)e.dst);
            src.push_back((int32_t)e.src);
            code.push_back((uint8_t)e.code);
        }
    ctx->n_local = (int64_t)dst.size();
    int rc;
    if ((rc = upload(ctx, &ctx->d_halo_dst, dst))) return rc;
    if ((rc = upload(ctx, &ctx->d_halo_src, src))) return rc;
    if ((rc = upload(ctx, &ctx->d_halo_code, code))) return rc;

    // push descriptors: where does the interior boundary strip next to (tile, edge) go?
    const int p = ctx->p;
    std::vector<int4> push((size_t)ctx->T * 4);
    ctx->push_complete = true;
    for (int k = 0; k < ctx->T; ++k) {
        const int t = ctx->order[ctx->t0 + k];
        const int face = t / (p * p), ty = (t % (p * p)) / p, tx = t % p;
        for (int e = 0; e < 4; ++e) {
            const bool inside = e == 0 ? ty < p - 1 : e == 1 ? ty > 0 : e == 2 ? tx < p - 1 : tx > 0;
            int dt_, de, rev = 0, code_ = 0;
            if (inside) {
                int dty = ty + (e == 0) - (e == 1), dtx = tx + (e == 2) - (e == 3);
                dt_ = face * p * p + dty * p + dtx;
                de = e ^ 1;   // N<->S, E<->W
            } else {
                SeamInfo s = cs_seam(face, e);
                SeamInfo back = cs_seam(s.fb, s.eb);   // our components seen from the neighbour
                int k = e < 2 ? tx : ty;
                if (s.rev) k = p - 1 - k;
                int dty = s.eb == 0 ? p - 1 : s.eb == 1 ? 0 : k;
                int dtx = s.eb == 2 ? p - 1 : s.eb == 3 ? 0 : k;
                dt_ = s.fb * p * p + dty * p + dtx;
                de = s.eb;
                rev = s.rev;
                code_ = back.code;
            }
            const int li = ctx->local_index(dt_);
            if (li < 0) ctx->push_complete = false;
            push[(size_t)k * 4 + e] = make_int4(li, de, rev, code_);
        }
    }
    return upload(ctx, &ctx->d_push, push);
}

int cs_build_peer_plan(cs_ctx* ctx, int n_peers, const int32_t* tb, const int32_t* tc) {
    if (n_peers < 1 || !tb || !tc) return cs_fail(ctx, CS_E_INVALID, "cs_halo_plan: need at least one peer");
    // peers are ranges of POSITIONS along axis 0; a tile belongs to the peer whose range holds its position
    auto owner = [&](int tile) {
        const int pos = ctx->inv[tile];
        for (int r = 0; r < n_peers; ++r)
            if (pos >= tb[r] && pos < tb[r] + tc[r]) return r;
        return -1;
    };
    for (int r = 0; r < n_peers; ++r)
        if (tb[r] < 0 || tc[r] < 0 || tb[r] + tc[r] > ctx->T_all)
            return cs_fail(ctx, CS_E_INVALID, "cs_halo_plan: peer %d owns positions [%d, %d), outside [0, %d)", r, tb[r], tb[r] + tc[r], ctx->T_all);
    ctx->peers.assign(n_peers, PeerPlan());
    std::vector<std::vector<int32_t>> send_src(n_peers), recv_dst(n_peers);
    std::vector<std::vector<uint8_t>> recv_code(n_peers);
    const int64_t tile_elems = (int64_t)ctx->m * ctx->m;
    // what we receive: our ghost cells whose source tile is owned by peer r, in table order
    for (const HaloEntry& e : ctx->halo) {
        if (e.src >= 0) continue;
        int r = owner(e.src_tile);
        if (r < 0) return cs_fail(ctx, CS_E_INVALID, "cs_halo_plan: tile %d is owned by no peer", e.src_tile);
        recv_dst[r].push_back((int32_t)e.dst);
        recv_code[r].push_back((uint8_t)e.code);
    }
    // what we send: every peer's ghost cells whose source is one of our tiles, in the peer's table order
    for (int r = 0; r < n_peers; ++r) {
        ctx->peers[r].tile_begin = tb[r];
        ctx->peers[r].tile_count = tc[r];
        if (tb[r] == ctx->t0 && tc[r] == ctx->T) continue;   // ourselves
        for (int pos = tb[r]; pos < tb[r] + tc[r]; ++pos)
            for_each_ghost(ctx->order[pos], ctx->N, ctx->p, ctx->g, [&](int, int, int ts, int32_t src_local, int) {
                const int li = ctx->local_index(ts);
                if (li >= 0 && owner(ts) != r) send_src[r].push_back((int32_t)((int64_t)li * tile_elems + src_local));
            });
    }
    std::vector<int32_t> all_send, all_recv;
    std::vector<uint8_t> all_code;
    int64_t so = 0, ro = 0;
    for (int r = 0; r < n_peers; ++r) {
        PeerPlan& pp = ctx->peers[r];
        pp.send_count = (int64_t)send_src[r].size();
        pp.recv_count = (int64_t)recv_dst[r].size();
        pp.send_offset = so;
        pp.recv_offset = ro;
        so += pp.send_count;
        ro += pp.recv_count;
        all_send.insert(all_send.end(), send_src[r].begin(), send_src[r].end());
        all_recv.insert(all_recv.end(), recv_dst[r].begin(), recv_dst[r].end());
        all_code.insert(all_code.end(), recv_code[r].begin(), recv_code[r].end());
    }
    ctx->send_total = so;
    ctx->recv_total = ro;
    int rc;
    if ((rc = upload(ctx, &ctx->d_send_src, all_send))) return rc;
    if ((rc = upload(ctx, &ctx->d_recv_dst, all_recv))) return rc;
    return upload(ctx, &ctx->d_recv_code, all_code);
}
