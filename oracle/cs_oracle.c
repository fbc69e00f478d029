/*
 * TEST INFRASTRUCTURE -- plain-C restatement of the CPU oracle (oracle/cs_oracle_np.py),
 * used (a) by tests/ as a second, faster checker and (b) by bench.py as the timed CPU
 * baseline / reference arm (OpenMP over faces and rows, all host cores).
 *
 * PARITY UNPINNED UPSTREAM: the reference ships no runnable solver, tests or golden vectors
 * (see the header of cs_oracle_np.py); this file is pinned to the NumPy oracle by
 * tests/test_oracle.py (agreement to round-off on every case) and shares its conventions:
 *   - layout field[face][j][i] with a ghost ring of width g, ghosts copied from the
 *     neighbour's interior (reference JAX-DevLab-Examples.py:141,174,183-195);
 *   - 'R' reverses the along-edge index, 'T' swaps row-strips and column-strips
 *     (JAX-DevLab-Examples.py:143-163);
 *   - numerics: DESIGN.md section 3 (clean-room; the reference has none).
 * The product never links or loads this file.
 *
 * Build: make -C oracle   (gcc -O3 -fopenmp -ffp-contract=off; no FMA contraction so that the
 * results track the NumPy oracle to the last bits).  -DCSO_REAL=float builds the same source with
 * the state, the metric tables and the per-cell arithmetic in single precision (set-up and the
 * conservation sums stay in double): the CPU arm of bench.py in the GPU arm's precision.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#ifdef _OPENMP
#include <omp.h>
#endif

#ifndef CSO_REAL
#define CSO_REAL double
#endif
typedef CSO_REAL real;
#define R_(x) ((real)(x))
static inline real r_min(real a, real b) { return sizeof(real) == 8 ? (real)fmin(a, b) : (real)fminf((float)a, (float)b); }
static inline real r_max(real a, real b) { return sizeof(real) == 8 ? (real)fmax(a, b) : (real)fmaxf((float)a, (float)b); }
static inline real r_abs(real a) { return sizeof(real) == 8 ? (real)fabs(a) : (real)fabsf((float)a); }

int cso_real_bytes(void) { return (int)sizeof(real); }

typedef struct {
    int N, g, M;
    int limiter, rk, advect_only;
    double a, omega, grav, delta;
    real *sg, *isg, *g11, *g12, *g22, *gi11, *gi12, *gi22;     /* [M*M] */
    real *fsg;                                                 /* [6*M*M] */
    real *hs;                                                  /* [6*M*M] */
    real *w[14];                                               /* work arrays [6*M*M] */
} cso_ctx;

static const int FR[6][3][3] = {
    {{0, 0, 1}, {1, 0, 0}, {0, 1, 0}},   {{0, 1, 0}, {-1, 0, 0}, {0, 0, 1}},
    {{-1, 0, 0}, {0, -1, 0}, {0, 0, 1}}, {{0, -1, 0}, {1, 0, 0}, {0, 0, 1}},
    {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}},   {{0, 0, -1}, {-1, 0, 0}, {0, 1, 0}},
};

static int dot3(const int* a, const int* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }

/* e: 0 N, 1 S, 2 E, 3 W */
static void edge_vec(int f, int e, int* o, int* t) {
    for (int k = 0; k < 3; ++k) {
        int x = FR[f][1][k], y = FR[f][2][k];
        o[k] = e == 0 ? y : e == 1 ? -y : e == 2 ? x : -x;
        t[k] = e < 2 ? x : y;
    }
}

typedef struct { int fb, eb, rev, S[2][2]; } seam_t;

static seam_t seam(int f, int e) {
    seam_t s;
    int oa[3], ta[3], ob[3], tb[3];
    edge_vec(f, e, oa, ta);
    s.fb = s.eb = -1;
    for (int k = 0; k < 6; ++k)
        if (FR[k][0][0] == oa[0] && FR[k][0][1] == oa[1] && FR[k][0][2] == oa[2]) s.fb = k;
    for (int k = 0; k < 4; ++k) {
        edge_vec(s.fb, k, ob, tb);
        if (ob[0] == FR[f][0][0] && ob[1] == FR[f][0][1] && ob[2] == FR[f][0][2]) { s.eb = k; break; }
    }
    edge_vec(s.fb, s.eb, ob, tb);
    s.rev = dot3(ta, tb) < 0;
    for (int col = 0; col < 2; ++col) {
        const int* w = FR[s.fb][1 + col];
        int img[3];
        for (int k = 0; k < 3; ++k) img[k] = dot3(w, ta) * ta[k] - dot3(w, ob) * oa[k];
        s.S[0][col] = dot3(img, FR[f][1]);
        s.S[1][col] = dot3(img, FR[f][2]);
    }
    return s;
}

/* Fill the ghosts of h (scalar, may be NULL) and (u, v) (vector pair, may be NULL). */
void cso_halo_fill(int N, int g, real* h, real* u, real* v) {
    const int M = N + 2 * g;
    const size_t FS = (size_t)M * M;
    for (int f = 0; f < 6; ++f)
        for (int e = 0; e < 4; ++e) {
            seam_t s = seam(f, e);
            for (int d = 0; d < g; ++d)
                for (int k = 0; k < N; ++k) {
                    int kk = s.rev ? N - 1 - k : k;
                    int jd, id, js, is;
                    switch (e) {
                        case 0: jd = g + N + d, id = g + k; break;
                        case 1: jd = g - 1 - d, id = g + k; break;
                        case 2: jd = g + k, id = g + N + d; break;
                        default: jd = g + k, id = g - 1 - d; break;
                    }
                    switch (s.eb) {
                        case 0: js = g + N - 1 - d, is = g + kk; break;
                        case 1: js = g + d, is = g + kk; break;
                        case 2: js = g + kk, is = g + N - 1 - d; break;
                        default: js = g + kk, is = g + d; break;
                    }
                    size_t od = f * FS + (size_t)jd * M + id, os = s.fb * FS + (size_t)js * M + is;
                    if (h) h[od] = h[os];
                    if (u) {
                        real src[2] = {u[os], v[os]}, out[2];
                        for (int r = 0; r < 2; ++r) {
                            int col = s.S[r][0] != 0 ? 0 : 1;
                            out[r] = s.S[r][col] < 0 ? -src[col] : src[col];
                        }
                        u[od] = out[0], v[od] = out[1];
                    }
                }
        }
    /* cube-vertex blocks: constant extrapolation of the E/W ghost columns in y */
    real* fields[3] = {h, u, v};
    for (int q = 0; q < 3; ++q) {
        real* a = fields[q];
        if (!a) continue;
        for (int f = 0; f < 6; ++f)
            for (int d = 0; d < g; ++d)
                for (int c = 0; c < g; ++c) {
                    int cols[2] = {c, g + N + c};
                    for (int s2 = 0; s2 < 2; ++s2) {
                        a[f * FS + (size_t)d * M + cols[s2]] = a[f * FS + (size_t)g * M + cols[s2]];
                        a[f * FS + (size_t)(g + N + d) * M + cols[s2]] = a[f * FS + (size_t)(g + N - 1) * M + cols[s2]];
                    }
                }
    }
}

cso_ctx* cso_create(int N, int g, int limiter, int rk, int advect_only, double a, double omega, double grav,
                    const double* axis, const double* tangents /* N */, const real* hs /* [6][M][M] or NULL */) {
    cso_ctx* c = (cso_ctx*)calloc(1, sizeof(cso_ctx));
    const int M = N + 2 * g;
    const size_t FS = (size_t)M * M;
    c->N = N, c->g = g, c->M = M, c->limiter = limiter, c->rk = rk, c->advect_only = advect_only;
    c->a = a, c->omega = omega, c->grav = grav, c->delta = 0.5 * 3.14159265358979323846 / N;
    real** tabs[] = {&c->sg, &c->isg, &c->g11, &c->g12, &c->g22, &c->gi11, &c->gi12, &c->gi22};
    for (int k = 0; k < 8; ++k) *tabs[k] = (real*)malloc(FS * sizeof(real));
    c->fsg = (real*)malloc(6 * FS * sizeof(real));
    c->hs = (real*)calloc(6 * FS, sizeof(real));
    if (hs) memcpy(c->hs, hs, 6 * FS * sizeof(real));
    for (int k = 0; k < 14; ++k) c->w[k] = (real*)calloc(6 * FS, sizeof(real));
    double* xt = (double*)malloc(M * sizeof(double));
    for (int I = 0; I < M; ++I) xt[I] = tangents[((I - g) % N + N) % N];
    const double ia2 = 1.0 / (a * a);
    for (int j = 0; j < M; ++j)
        for (int i = 0; i < M; ++i) {
            const double X = xt[i], Y = xt[j];
            const double A = 1.0 + X * X, B = 1.0 + Y * Y, rho2 = A + Y * Y, r2inv = 1.0 / rho2;
            const double cc = (a * a) * A * B * (r2inv * r2inv);
            const size_t o = (size_t)j * M + i;
            const double sg = cc * sqrt(rho2);            /* tables are computed in double and stored as real */
            c->sg[o] = R_(sg);
            c->isg[o] = R_(1.0 / sg);
            c->g11[o] = R_(cc * A), c->g12[o] = R_(-(cc * X) * Y), c->g22[o] = R_(cc * B);
            c->gi11[o] = R_((rho2 * ia2) / A);
            c->gi12[o] = R_(((rho2 * ia2) * (X * Y)) / (A * B));
            c->gi22[o] = R_((rho2 * ia2) / B);
            for (int f = 0; f < 6; ++f) {
                double cf[3];
                for (int k = 0; k < 3; ++k) cf[k] = axis[0] * FR[f][k][0] + axis[1] * FR[f][k][1] + axis[2] * FR[f][k][2];
                c->fsg[f * FS + o] = R_((2.0 * omega) * (cf[0] + cf[1] * X + cf[2] * Y) * cc);
            }
        }
    free(xt);
    return c;
}

void cso_destroy(cso_ctx* c) {
    if (!c) return;
    real* p[] = {c->sg, c->isg, c->g11, c->g12, c->g22, c->gi11, c->gi12, c->gi22, c->fsg, c->hs};
    for (int k = 0; k < 10; ++k) free(p[k]);
    for (int k = 0; k < 14; ++k) free(c->w[k]);
    free(c);
}

static inline real sgn(real x) { return R_((x > R_(0)) - (x < R_(0))); }

static inline real slope(real dm, real dp, int lim) {
    if (lim == 0) return R_(0.5) * (dm + dp);
    const real s = R_(0.5) * (sgn(dm) + sgn(dp));
    if (lim == 1) return s * r_min(r_abs(dm), r_abs(dp));
    return s * r_min(r_min(R_(2) * r_abs(dm), R_(2) * r_abs(dp)), R_(0.5) * r_abs(dm + dp));
}

/* out = ca * q0 + cb * (qs + dt * L(qs)) on interior cells; qs halo-filled */
static void stage(cso_ctx* c, const real* h0, const real* u0, const real* v0, const real* h, const real* u,
                  const real* v, real* ho, real* uo, real* vo, double ca_, double cb_, double dt_) {
    const int N = c->N, g = c->g, M = c->M, lim = c->limiter;
    const size_t FS = (size_t)M * M;
    real *m1 = c->w[6], *m2 = c->w[7], *c1 = c->w[8], *c2 = c->w[9], *Bn = c->w[10], *sx = c->w[11], *sy = c->w[12];
    const real rd = R_(1.0 / c->delta), r2d = R_(0.5 / c->delta), grav = R_(c->grav);
    /* the convex pair stays convex after rounding to real (a + b == 1 exactly), as in the product */
    const real cb = R_(cb_), ca = ca_ != 0.0 ? R_(1) - cb : R_(0), dt = R_(dt_);
#pragma omp parallel for collapse(2) schedule(static)
    for (int f = 0; f < 6; ++f)
        for (int j = 0; j < M; ++j) {
            const size_t row = f * FS + (size_t)j * M;
            const size_t trow = (size_t)j * M;
            for (int i = 0; i < M; ++i) {
                const size_t o = row + i, t = trow + i;
                const real uu = u[o], vv = v[o];
                m1[o] = c->sg[t] * uu;
                m2[o] = c->sg[t] * vv;
                const real a1 = c->g11[t] * uu + c->g12[t] * vv, a2 = c->g12[t] * uu + c->g22[t] * vv;
                c1[o] = a1, c2[o] = a2;
                Bn[o] = grav * (h[o] + c->hs[o]) + R_(0.5) * (uu * a1 + vv * a2);
                sx[o] = (i >= 1 && i < M - 1) ? slope(h[o] - h[o - 1], h[o + 1] - h[o], lim) : R_(0);
                sy[o] = (j >= 1 && j < M - 1) ? slope(h[o] - h[o - M], h[o + M] - h[o], lim) : R_(0);
            }
        }
#pragma omp parallel for collapse(2) schedule(static)
    for (int f = 0; f < 6; ++f)
        for (int j = g; j < g + N; ++j) {
            const size_t row = f * FS + (size_t)j * M;
            for (int i = g; i < g + N; ++i) {
                const size_t o = row + i, t = (size_t)j * M + i;
                const real Z = R_(0), H = R_(0.5);
                real Fm, FE, FW, GN, GS;
                Fm = H * (m1[o] + m1[o + 1]);
                FE = r_max(Fm, Z) * (h[o] + H * sx[o]) + r_min(Fm, Z) * (h[o + 1] - H * sx[o + 1]);
                Fm = H * (m1[o - 1] + m1[o]);
                FW = r_max(Fm, Z) * (h[o - 1] + H * sx[o - 1]) + r_min(Fm, Z) * (h[o] - H * sx[o]);
                Fm = H * (m2[o] + m2[o + M]);
                GN = r_max(Fm, Z) * (h[o] + H * sy[o]) + r_min(Fm, Z) * (h[o + M] - H * sy[o + M]);
                Fm = H * (m2[o - M] + m2[o]);
                GS = r_max(Fm, Z) * (h[o - M] + H * sy[o - M]) + r_min(Fm, Z) * (h[o] - H * sy[o]);
                const real dh = -(((FE - FW) + (GN - GS)) * rd) * (R_(1) / c->sg[t]);
                real du = Z, dv = Z;
                if (!c->advect_only) {
                    const real eta = ((c2[o + 1] - c2[o - 1]) - (c1[o + M] - c1[o - M])) * r2d + c->fsg[f * FS + t];
                    const real T1 = eta * v[o] - (Bn[o + 1] - Bn[o - 1]) * r2d;
                    const real T2 = -eta * u[o] - (Bn[o + M] - Bn[o - M]) * r2d;
                    du = c->gi11[t] * T1 + c->gi12[t] * T2;
                    dv = c->gi12[t] * T1 + c->gi22[t] * T2;
                }
                ho[o] = ca * h0[o] + cb * (h[o] + dt * dh);
                uo[o] = ca * u0[o] + cb * (u[o] + dt * du);
                vo[o] = ca * v0[o] + cb * (v[o] + dt * dv);
            }
        }
    cso_halo_fill(N, g, ho, uo, vo);
}

/* nsteps explicit steps in place; h, u, v halo-filled [6][M][M] */
void cso_step(cso_ctx* c, real* h, real* u, real* v, double dt, int nsteps) {
    static const double PA[3][3] = {{0, 0, 0}, {0, 0.5, 0}, {0, 0.75, 0}};
    const int ns = c->rk;   /* 1 euler, 2 ssprk2, 3 ssprk3 */
    double ca[3] = {0.0, PA[ns - 1][1], 1.0 - 2.0 / 3.0}, cb[3] = {1.0, 1.0 - PA[ns - 1][1], 2.0 / 3.0};
    if (ns == 2) ca[1] = 0.5, cb[1] = 0.5;
    if (ns == 3) ca[1] = 0.75, cb[1] = 0.25;
    const size_t bytes = 6 * (size_t)c->M * c->M * sizeof(real);
    for (int it = 0; it < nsteps; ++it) {
        real *qa[3] = {c->w[0], c->w[1], c->w[2]}, *qb[3] = {c->w[3], c->w[4], c->w[5]};
        const real *ch = h, *cu = u, *cv = v;
        for (int s = 0; s < ns; ++s) {
            real** dst = (s % 2 == 0) ? qa : qb;
            stage(c, h, u, v, ch, cu, cv, dst[0], dst[1], dst[2], ca[s], cb[s], dt);
            ch = dst[0], cu = dst[1], cv = dst[2];
        }
        memcpy(h, ch, bytes), memcpy(u, cu, bytes), memcpy(v, cv, bytes);
    }
}

void cso_diagnostics(cso_ctx* c, const real* h, const real* u, const real* v, double* out) {
    const int N = c->N, g = c->g, M = c->M;
    const size_t FS = (size_t)M * M;
    double mass = 0.0, energy = 0.0;
    const double da = c->delta * c->delta;
#pragma omp parallel for collapse(2) reduction(+ : mass, energy) schedule(static)
    for (int f = 0; f < 6; ++f)
        for (int j = g; j < g + N; ++j)
            for (int i = g; i < g + N; ++i) {
                const size_t o = f * FS + (size_t)j * M + i, t = (size_t)j * M + i;
                /* always in double, whatever real is */
                const double hh = h[o], uu = u[o], vv = v[o], hs = c->hs[o];
                const double a1 = (double)c->g11[t] * uu + (double)c->g12[t] * vv, a2 = (double)c->g12[t] * uu + (double)c->g22[t] * vv;
                const double K = 0.5 * (uu * a1 + vv * a2), w = (double)c->sg[t] * da;
                mass += w * hh;
                energy += w * (hh * K + 0.5 * c->grav * hh * hh + c->grav * hh * hs);
            }
    out[0] = mass, out[1] = energy;
}

/* torchrun exports OMP_NUM_THREADS=1 to its workers; the timed baseline asks for the cores it may use */
void cso_set_num_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

int cso_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
