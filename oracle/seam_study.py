"""TEST INFRASTRUCTURE -- study, not part of the product or of the parity oracle.

Order of accuracy at the 12 panel seams for three ways of building the ghost cells, measured with the NumPy oracle on
Williamson case 2 (steady geostrophic flow; the exact solution is the initial state):

  copy      ghost = the neighbour's own cell (same along-edge index), vector components through the seam's signed
            permutation, metric continued periodically.  The reference's halo semantics (JAX-DevLab-Examples.py:
            183-195: "write to ghost cells" the neighbour's boundary data) and the product's default; every seam flux
            is single valued, so mass is conserved to round-off.
  rotate    the same copy, but vectors re-expressed in OUR extended chart by the geometric 2 x 2 matrix at the
            neighbour's cell centre (product: geometry.seam_rotation_tables), metric = our chart
            extended past +-pi/4.
  interp    ghost values at OUR extended grid points: 1-D cubic interpolation along the neighbour's cell row at the
            same depth (Putman & Lin 2007; on the equiangular grid the ghost points lie exactly on the neighbour's
            coordinate lines), vectors interpolated then rotated at the ghost point, metric extended.

Run:  python -m oracle.seam_study [--N 24 48 96] [--days 1.0]
Prints the normalised L2 error of h, the observed order between resolutions and the relative mass drift.
Written to answer VERDICT r1 item 6; the table it prints is recorded in DESIGN.md section 3.5.
"""
from __future__ import annotations

import argparse
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
for q in (str(ROOT / "sharding-the-sphere-fall-2025-jax-devlab-examples_b200"), str(ROOT)):
    if q not in sys.path:
        sys.path.insert(0, q)

from oracle import cs_oracle_np as orc  # noqa: E402

FR = orc.FRAMES.astype(np.float64)


def _basis(face, P, a):
    """d r/d alpha, d r/d beta of `face`'s chart at unit points P [K, 3] (face: int array [K])."""
    nrm, ex, ey = FR[face, 0], FR[face, 1], FR[face, 2]
    pn = np.sum(P * nrm, -1)
    X, Y = np.sum(P * ex, -1) / pn, np.sum(P * ey, -1) / pn
    rho = np.sqrt(1.0 + X * X + Y * Y)
    Q = nrm + X[:, None] * ex + Y[:, None] * ey
    ea = a * (1.0 + X * X)[:, None] * (ex / rho[:, None] - Q * (X / rho ** 3)[:, None])
    eb = a * (1.0 + Y * Y)[:, None] * (ey / rho[:, None] - Q * (Y / rho ** 3)[:, None])
    return ea, eb, X, Y


def _rotation(dst_face, src_face, P, a):
    """Contravariant components in src's chart -> contravariant components in dst's chart, at points P."""
    sa, sb, _, _ = _basis(src_face, P, a)
    da, db, _, _ = _basis(dst_face, P, a)
    Bs, Bd = np.stack([sa, sb], -1), np.stack([da, db], -1)
    G = np.einsum("kxi,kxj->kij", Bd, Bd)
    M = np.einsum("kxi,kxj->kij", Bd, Bs)
    return np.linalg.solve(G, M)


class SeamHalo:
    """Table-driven ghost fill for the 'rotate' and 'interp' constructions (edge strips only: the stencil reads no
    corner ghosts).  Each ghost cell is sum_q w[q] * src[q] over (up to) four cells of one neighbour row, then R."""

    def __init__(self, N, g, mode, a=orc.EARTH_RADIUS, points=4):
        self.N, self.g, self.mode = N, g, mode
        M = N + 2 * g
        d = 0.5 * np.pi / N
        ang = -0.25 * np.pi + (np.arange(-g, N + g) + 0.5) * d
        self.tangents = np.tan(ang)                        # our chart, extended past the edges
        J, I = np.meshgrid(np.arange(M), np.arange(M), indexing="ij")
        inj, ini = (J >= g) & (J < g + N), (I >= g) & (I < g + N)
        strip = inj ^ ini                                  # exactly one index outside: edge strips, no corner blocks
        gj, gi = J[strip], I[strip]
        dst, src, wts, rot = [], [], [], []
        for L in range(6):
            X, Y = self.tangents[gi], self.tangents[gj]
            if mode == "rotate":
                # the copy's source: same along-edge index, neighbour's interior at the same depth; its own centre
                fcopy = np.zeros((6, M, M))
                fcopy[:, g:g + N, g:g + N] = np.arange(6 * N * N, dtype=np.float64).reshape(6, N, N)
                filled = orc.halo_fill_scalar(fcopy, g)[L][gj, gi].astype(np.int64)
                B, rem = filled // (N * N), filled % (N * N)
                jb, ib = rem // N, rem % N
                t = orc.centre_tangents(N)
                Ps = FR[B, 0] + t[ib][:, None] * FR[B, 1] + t[jb][:, None] * FR[B, 2]
                Ps /= np.linalg.norm(Ps, axis=-1, keepdims=True)
                base = (B * M + jb + g) * M + ib + g
                idx = np.stack([base] * points, -1)
                w = np.zeros((gj.size, points))
                w[:, 0] = 1.0
                R = _rotation(np.full(gj.size, L), B, Ps, a)
            else:
                P = FR[L, 0] + X[:, None] * FR[L, 1] + Y[:, None] * FR[L, 2]
                P /= np.linalg.norm(P, axis=-1, keepdims=True)
                B = np.argmax(P @ FR[:, 0].T, axis=1)                     # the face the ghost point lies on
                _, _, XB, YB = _basis(B, P, a)
                fi = (np.arctan(XB) + 0.25 * np.pi) / d - 0.5             # fractional interior indices in B
                fj = (np.arctan(YB) + 0.25 * np.pi) / d - 0.5
                ri, rj = np.abs(fi - np.round(fi)), np.abs(fj - np.round(fj))
                assert np.all(np.minimum(ri, rj) < 1e-6), "a ghost point is not on a coordinate line of its neighbour"
                along_j = ri <= rj                                         # i is the (integer) depth index: interpolate along j
                f = np.where(along_j, fj, fi)
                fixed = np.round(np.where(along_j, fi, fj)).astype(np.int64)
                assert np.all((fixed >= 0) & (fixed < N))
                b0 = np.clip(np.floor(f).astype(np.int64) - (points // 2 - 1), 0, N - points)
                nodes = b0[:, None] + np.arange(points)[None, :]
                w = np.ones((gj.size, points))
                for q in range(points):                                   # Lagrange weights on the uniform angle grid
                    for r in range(points):
                        if r != q:
                            w[:, q] *= (f - nodes[:, r]) / (nodes[:, q] - nodes[:, r])
                jb = np.where(along_j[:, None], nodes, fixed[:, None])
                ib = np.where(along_j[:, None], fixed[:, None], nodes)
                idx = (B[:, None] * M + jb + g) * M + ib + g
                R = _rotation(np.full(gj.size, L), B, P, a)
            dst.append((L * M + gj) * M + gi)
            src.append(idx)
            wts.append(w)
            rot.append(R)
        self.dst, self.src = np.concatenate(dst), np.concatenate(src)
        self.w, self.R = np.concatenate(wts), np.concatenate(rot)

    def __call__(self, h, u, v):
        hf, uf, vf = h.reshape(-1), u.reshape(-1), v.reshape(-1)
        hf[self.dst] = np.sum(self.w * hf[self.src], -1)
        a = np.sum(self.w * uf[self.src], -1)
        b = np.sum(self.w * vf[self.src], -1)
        uf[self.dst] = self.R[:, 0, 0] * a + self.R[:, 0, 1] * b
        vf[self.dst] = self.R[:, 1, 0] * a + self.R[:, 1, 1] * b
        return h, u, v


def run_case(N, mode, days, alpha0=0.0, limiter="none", case_name="tc2"):
    from cubesphere import cases            # initial conditions (host-side set-up code, not a kernel)
    g = 2
    case = cases.make_case(case_name, N, alpha0=alpha0) if case_name == "tc2" else cases.make_case(case_name, N)
    if mode == "copy":
        p = orc.Params(N, g, axis=case.axis, limiter=limiter, hs=case.hs)
        st = orc.halo_fill_state(orc.pad(case.h, g), orc.pad(case.u1, g), orc.pad(case.u2, g), g)
    else:
        halo = SeamHalo(N, g, mode)
        p = orc.Params(N, g, axis=case.axis, limiter=limiter, hs=None, tangents=halo.tangents, halo_fill=halo)
        if case.hs is not None:
            hs = orc.pad(case.hs, g)
            halo(hs, np.zeros_like(hs), np.zeros_like(hs))
            p.hs = hs
        st = halo(orc.pad(case.h, g), orc.pad(case.u1, g), orc.pad(case.u2, g))
    dt = cases.stable_dt(case, 0.5)
    nsteps = int(round(days * 86400.0 / dt))
    c = slice(g, g + N)
    h0 = st[0].copy()
    m0, _ = orc.diagnostics(p, *st)
    for _ in range(nsteps):
        st = orc.step(p, st, dt)
    m1, _ = orc.diagnostics(p, *st)
    w = p.sg[c, c]
    err = np.sqrt(np.sum(w * (st[0][:, c, c] - h0[:, c, c]) ** 2) / np.sum(w * h0[:, c, c] ** 2))
    return float(err), float((m1 - m0) / m0), nsteps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--N", type=int, nargs="+", default=[24, 48, 96])
    ap.add_argument("--days", type=float, default=1.0)
    ap.add_argument("--alpha0", type=float, default=0.0)
    ap.add_argument("--modes", nargs="+", default=["copy", "rotate", "interp"])
    args = ap.parse_args()
    print(f"Williamson case 2, alpha0 = {args.alpha0}, {args.days} day(s), limiter none, SSP-RK3, CFL 0.5")
    print(f"{'mode':8s} " + " ".join(f"{'C%d' % n:>12s}" for n in args.N) + "   orders        mass drift (finest)")
    for mode in args.modes:
        errs, drift = [], 0.0
        for n in args.N:
            e, drift, _ = run_case(n, mode, args.days, args.alpha0)
            errs.append(e)
        orders = [np.log2(errs[k] / errs[k + 1]) / np.log2(args.N[k + 1] / args.N[k]) for k in range(len(errs) - 1)]
        print(f"{mode:8s} " + " ".join(f"{e:12.4e}" for e in errs) + "   " + " ".join(f"{o:5.2f}" for o in orders) + f"   {drift:+.2e}")


if __name__ == "__main__":
    main()
