"""Equiangular gnomonic cubed-sphere geometry (host side, fp64).

Not in the reference (its "Geometry: Math/Mesh" box, deck slide 6, is not
shipped); written from the published equiangular metric (Nair, Thomas & Loft
2005; Ullrich, Jablonowski & van Leer 2010).

Point on face f:  r = a * (n_f + X x_f + Y y_f) / rho,  X = tan(alpha),
Y = tan(beta), rho^2 = 1 + X^2 + Y^2, alpha, beta in [-pi/4, pi/4].

The metric is the same on every panel and depends on (X, Y) only, so the
kernels need one 1-D table ``xtab`` of X at cell centres.  Ghost entries are the
*periodic* continuation ``xtab[g + N + d] = xtab[g + d]``: a ghost cell holds a
copy of the neighbour's cell, and seen in the unfolded frame that cell's own
chart coordinates are exactly the periodic continuation (see DESIGN.md,
"ghost-cell metric").  All ghost-cell quantities the stencil uses (sqrt(G) u^n,
kinetic energy, potential) are then the neighbour's own values, which makes
every seam flux single valued and mass conservation exact.
"""
from __future__ import annotations

import numpy as np

from .topology import FACE_FRAMES

EARTH_RADIUS = 6.37122e6      # m
EARTH_OMEGA = 7.292e-5        # 1/s
EARTH_GRAVITY = 9.80616       # m/s^2


def centre_tangents(N: int) -> np.ndarray:
    """X_i = tan(alpha_i), alpha_i = -pi/4 + (i + 1/2) pi/(2N); exactly antisymmetric."""
    i = np.arange(N, dtype=np.float64)
    x = np.tan(-0.25 * np.pi + (i + 0.5) * (0.5 * np.pi / N))
    half = N // 2
    x[:half] = -x[::-1][:half]
    if N % 2:
        x[half] = 0.0
    return x


def padded_tangents(N: int, g: int, lo: int = 0, n: int | None = None) -> np.ndarray:
    """Periodic continuation of :func:`centre_tangents` for the tile starting at
    face-level index ``lo`` with ``n`` interior cells (whole face by default)."""
    n = N if n is None else n
    x = centre_tangents(N)
    idx = (lo + np.arange(-g, n + g)) % N
    return x[idx]


def metric_terms(X, Y, a=EARTH_RADIUS):
    """Point-wise metric terms from broadcastable X, Y (the oracle's definition).

    Returns dict with sg = sqrt(G), covariant g11, g12, g22 and contravariant
    gi11, gi12, gi22, plus c = a^2 (1+X^2)(1+Y^2) / rho^4 (so f sqrt(G) = f_num c).
    """
    A = 1.0 + X * X
    B = 1.0 + Y * Y
    rho2 = A + Y * Y
    r2inv = 1.0 / rho2
    c = (a * a) * A * B * (r2inv * r2inv)
    sg = c * np.sqrt(rho2)
    ia2 = 1.0 / (a * a)
    return dict(
        sg=sg, c=c, rho2=rho2,
        g11=c * A, g12=-(c * X) * Y, g22=c * B,
        gi11=(rho2 * ia2) / A, gi12=((rho2 * ia2) * (X * Y)) / (A * B), gi22=(rho2 * ia2) / B,
    )


def face_points(N: int, g: int = 0):
    """Unit position vectors of cell centres, shape [6, N+2g, N+2g, 3].

    Ghost entries (g > 0) use the *extended* chart of the face (alpha beyond
    +-pi/4), which is only meant for plotting/diagnostics; initial conditions are
    sampled at interior cells and ghosts are filled by the halo exchange.
    """
    i = np.arange(-g, N + g, dtype=np.float64)
    t = np.tan(-0.25 * np.pi + (i + 0.5) * (0.5 * np.pi / N))
    if g == 0:
        t = centre_tangents(N)
    X, Y = np.meshgrid(t, t, indexing="xy")      # X varies along axis 1 (i), Y along axis 0 (j)
    out = np.empty((6,) + X.shape + (3,), dtype=np.float64)
    for f, (nrm, ex, ey) in enumerate(FACE_FRAMES):
        nrm, ex, ey = (np.asarray(v, dtype=np.float64) for v in (nrm, ex, ey))
        r = nrm[None, None, :] + X[..., None] * ex + Y[..., None] * ey
        out[f] = r / np.linalg.norm(r, axis=-1, keepdims=True)
    return out


class Patch:
    """Cell-centre coordinates of a set of tiles: faces [K], X, Y [K, n, n] (X along axis 2, Y along axis 1).

    ``tiles=None`` means all 6 p^2 tiles in order; with p = 1 that is the six whole faces.  Initial
    conditions are evaluated patch by patch so that a rank only ever builds its own shard.
    """

    def __init__(self, N, p=1, tiles=None):
        if N % p:
            raise ValueError(f"N = {N} is not divisible by tiles_per_edge = {p}")
        n = N // p
        t = centre_tangents(N)
        tiles = list(range(6 * p * p)) if tiles is None else list(tiles)
        self.N, self.p, self.n, self.tiles = N, p, n, tiles
        self.faces = np.empty(len(tiles), dtype=np.int64)
        self.X = np.empty((len(tiles), n, n))
        self.Y = np.empty((len(tiles), n, n))
        for k, tile in enumerate(tiles):
            f, r = divmod(tile, p * p)
            ty, tx = divmod(r, p)
            self.faces[k] = f
            self.X[k], self.Y[k] = np.meshgrid(t[tx * n:(tx + 1) * n], t[ty * n:(ty + 1) * n], indexing="xy")

    def _frames(self):
        fr = np.asarray(FACE_FRAMES, dtype=np.float64)[self.faces]          # [K, 3(normal, x, y), 3]
        return fr[:, 0, None, None, :], fr[:, 1, None, None, :], fr[:, 2, None, None, :]

    def points(self):
        """Unit position vectors [K, n, n, 3]."""
        nrm, ex, ey = self._frames()
        r = nrm + self.X[..., None] * ex + self.Y[..., None] * ey
        return r / np.linalg.norm(r, axis=-1, keepdims=True)

    def basis(self, a=EARTH_RADIUS):
        """d r / d alpha, d r / d beta at the cell centres: two arrays [K, n, n, 3]."""
        nrm, ex, ey = self._frames()
        X, Y = self.X, self.Y
        rho = np.sqrt(1.0 + X * X + Y * Y)
        P = nrm + X[..., None] * ex + Y[..., None] * ey
        e_a = a * (1.0 + X * X)[..., None] * (ex / rho[..., None] - P * (X / rho ** 3)[..., None])
        e_b = a * (1.0 + Y * Y)[..., None] * (ey / rho[..., None] - P * (Y / rho ** 3)[..., None])
        return e_a, e_b


def lonlat(points):
    """Longitude in [0, 2 pi) and latitude from unit vectors [..., 3]."""
    lon = np.mod(np.arctan2(points[..., 1], points[..., 0]), 2.0 * np.pi)
    lat = np.arcsin(np.clip(points[..., 2], -1.0, 1.0))
    return lon, lat


def tangent_basis(N: int, a=EARTH_RADIUS):
    """d r / d alpha and d r / d beta at interior cell centres: two arrays [6, N, N, 3]."""
    t = centre_tangents(N)
    X, Y = np.meshgrid(t, t, indexing="xy")
    rho = np.sqrt(1.0 + X * X + Y * Y)
    e_a = np.empty((6, N, N, 3))
    e_b = np.empty((6, N, N, 3))
    for f, (nrm, ex, ey) in enumerate(FACE_FRAMES):
        nrm, ex, ey = (np.asarray(v, dtype=np.float64) for v in (nrm, ex, ey))
        P = nrm + X[..., None] * ex + Y[..., None] * ey            # un-normalised position
        e_a[f] = a * (1.0 + X * X)[..., None] * (ex / rho[..., None] - P * (X / rho ** 3)[..., None])
        e_b[f] = a * (1.0 + Y * Y)[..., None] * (ey / rho[..., None] - P * (Y / rho ** 3)[..., None])
    return e_a, e_b


def contravariant_from_cartesian(V, N: int, a=EARTH_RADIUS, patch=None):
    """Panel-local contravariant components (rad/s) of Cartesian velocities V [K, n, n, 3] (m/s)."""
    e_a, e_b = tangent_basis(N, a) if patch is None else patch.basis(a)
    g11 = np.sum(e_a * e_a, -1)
    g12 = np.sum(e_a * e_b, -1)
    g22 = np.sum(e_b * e_b, -1)
    b1 = np.sum(e_a * V, -1)          # covariant components
    b2 = np.sum(e_b * V, -1)
    det = g11 * g22 - g12 * g12
    return (g22 * b1 - g12 * b2) / det, (g11 * b2 - g12 * b1) / det


def cartesian_from_lonlat_wind(u_east, v_north, points):
    """Cartesian velocity from zonal/meridional components at unit points [..., 3]."""
    lon, lat = lonlat(points)
    e_lon = np.stack([-np.sin(lon), np.cos(lon), np.zeros_like(lon)], -1)
    e_lat = np.stack([-np.sin(lat) * np.cos(lon), -np.sin(lat) * np.sin(lon), np.cos(lat)], -1)
    return u_east[..., None] * e_lon + v_north[..., None] * e_lat


def coriolis_coefficients(axis=(0.0, 0.0, 1.0)) -> np.ndarray:
    """Per-face (c0, c1, c2) with sin(lat_rot) = (c0 + c1 X + c2 Y) / rho; shape [6, 3].

    ``axis`` is the unit rotation axis (tilted for the rotated Williamson cases).
    """
    axis = np.asarray(axis, dtype=np.float64)
    return np.array([[axis @ np.asarray(v, dtype=np.float64) for v in frame] for frame in FACE_FRAMES])


# ----------------------------------------------------------------------------
# vector halo: geometric seam rotation  (no counterpart in the reference, which exchanges
# Cartesian components instead -- deck slide 18)
# ----------------------------------------------------------------------------

def chart_basis(face, P, a=EARTH_RADIUS):
    """Tangent basis (d r / d alpha, d r / d beta) of ``face``'s gnomonic chart at unit points ``P [..., 3]``.

    The chart of a face covers its whole hemisphere, so ``P`` may lie beyond the face's own edges (ghost cells):
    X = (P . x_f) / (P . n_f), Y = (P . y_f) / (P . n_f).  ``face`` is an int or an int array broadcastable to
    ``P.shape[:-1]``.  Returns ``(e_a, e_b, X, Y)``.
    """
    fr = np.asarray(FACE_FRAMES, dtype=np.float64)[np.asarray(face)]
    nrm, ex, ey = fr[..., 0, :], fr[..., 1, :], fr[..., 2, :]
    pn = np.sum(P * nrm, -1)
    if np.any(pn <= 0):
        raise ValueError("a point lies outside the hemisphere of the chart it is to be expressed in")
    X = np.sum(P * ex, -1) / pn
    Y = np.sum(P * ey, -1) / pn
    rho = np.sqrt(1.0 + X * X + Y * Y)
    Q = nrm + X[..., None] * ex + Y[..., None] * ey
    e_a = a * (1.0 + X * X)[..., None] * (ex / rho[..., None] - Q * (X / rho ** 3)[..., None])
    e_b = a * (1.0 + Y * Y)[..., None] * (ey / rho[..., None] - Q * (Y / rho ** 3)[..., None])
    return e_a, e_b, X, Y


def seam_rotation(dst_face, src_face, src_X, src_Y, kind="contravariant", a=EARTH_RADIUS):
    """2 x 2 matrices that re-express a vector held in ``src_face``'s chart at the point (src_X, src_Y) of that chart
    in the (extended) chart of ``dst_face``:   (u, v)_dst = R (u, v)_src.

    With B = [d r/d alpha, d r/d beta] (3 x 2) of either chart at the same point P and G = B^T B:
      contravariant components (what this solver carries):  R = G_dst^-1 B_dst^T B_src,
      covariant components:                                 R = B_dst^T B_src G_src^-1.
    Same face -> the exact identity.  All arguments broadcast; the result has shape ``[..., 2, 2]``.
    """
    if kind not in ("contravariant", "covariant"):
        raise ValueError(f"kind must be 'contravariant' or 'covariant', got {kind!r}")
    dst_face, src_face = np.asarray(dst_face), np.asarray(src_face)
    src_X, src_Y = np.asarray(src_X, dtype=np.float64), np.asarray(src_Y, dtype=np.float64)
    fr = np.asarray(FACE_FRAMES, dtype=np.float64)[src_face]
    P = fr[..., 0, :] + src_X[..., None] * fr[..., 1, :] + src_Y[..., None] * fr[..., 2, :]
    P = P / np.linalg.norm(P, axis=-1, keepdims=True)
    sa, sb, _, _ = chart_basis(src_face, P, a)
    da, db, _, _ = chart_basis(np.broadcast_to(dst_face, src_X.shape), P, a)
    Bs = np.stack([sa, sb], -1)                           # [..., 3, 2]
    Bd = np.stack([da, db], -1)
    M = np.einsum("...ki,...kj->...ij", Bd, Bs)           # B_dst^T B_src
    if kind == "contravariant":
        G = np.einsum("...ki,...kj->...ij", Bd, Bd)
        R = np.linalg.solve(G, M)
    else:
        G = np.einsum("...ki,...kj->...ij", Bs, Bs)
        R = np.einsum("...ij,...jk->...ik", M, np.linalg.inv(G))
    same = np.broadcast_to(dst_face == src_face, R.shape[:-2])
    R = np.where(same[..., None, None], np.eye(2), R)
    return R


def seam_rotation_tables(N, g, p=1, kind="contravariant", a=EARTH_RADIUS, table=None):
    """Rotation matrix of every ghost cell, ``[n_ghost, 2, 2]`` fp64, host-precomputed.

    Ghost cell k of :func:`topology.build_halo_table` (or of ``table`` = a dict with the same keys, e.g. decoded from
    a context's ``halo_table()``) holds a copy of the interior cell ``(src_tile, src_j, src_i)``; row k re-expresses
    that cell's panel-local vector in the extended chart of the ghost's own face, at the source cell's centre.
    Corner blocks resolve to the E / W ghost column they extrapolate, so they get that column's matrix
    ("already-rotated components").  Pass the result to ``Context.set_rotation`` / ``cs_halo_set_rotation``.
    """
    from . import topology
    tab = topology.build_halo_table(N, p, g) if table is None else table
    n = N // p
    t = centre_tangents(N)
    pp = p * p
    src_face = tab["src_tile"] // pp
    dst_face = tab["dst_tile"] // pp
    r = tab["src_tile"] % pp
    sx = t[(r % p) * n + tab["src_i"] - g]
    sy = t[(r // p) * n + tab["src_j"] - g]
    return seam_rotation(dst_face, src_face, sx, sy, kind, a)


# ----------------------------------------------------------------------------
# second-order ghosts: values at OUR extended grid points  (option; the default halo is the reference's plain copy)
# ----------------------------------------------------------------------------

def extended_tangents(N: int, g: int) -> np.ndarray:
    """tan(alpha_I) for I = -g .. N+g-1 with the chart simply continued past +-pi/4 (N + 2g entries): the metric a
    ghost cell has when it is a point of OUR grid, as opposed to :func:`padded_tangents` (a copy of the neighbour's)."""
    x = np.tan(-0.25 * np.pi + (np.arange(-g, N + g) + 0.5) * (0.5 * np.pi / N))
    x[g:g + N] = centre_tangents(N)
    return x


def seam_interp_tables(N, g, p=1, points=4, a=EARTH_RADIUS, table=None):
    """Tables for ghost values at the points of our own extended grid (Putman & Lin 2007 style), which make the
    scheme second order at the seams where the plain copy is first order -- at the price of seam fluxes that are no
    longer single valued (mass drifts at truncation level).  See DESIGN.md 3.5 and oracle/seam_study.py.

    On the equiangular grid a ghost point at depth d lies exactly on the neighbour's coordinate line at depth d, so one
    1-D Lagrange interpolation (``points`` nodes, one-sided next to the cube corners so that only interior cells are
    read) along that line gives the value; vectors are interpolated component-wise in the neighbour's chart and then
    re-expressed in ours by the 2 x 2 matrix of :func:`seam_rotation` at the ghost point.

    Returns a dict, one row per ghost cell in the order of ``table`` (default :func:`topology.build_halo_table`):
      ``dst`` [n] element offsets, ``src`` [n, 4] element offsets, ``w`` [n, 4] weights, ``rot`` [n, 2, 2],
    plus ``tangents`` = :func:`extended_tangents`.  Ghost cells between tiles of one face and the corner blocks (which
    the stencil never reads) keep the plain copy: weight 1 on the copy's source, rot = the signed permutation.
    """
    from . import topology
    if points < 2 or points > 4:
        raise ValueError("points must be 2, 3 or 4")
    tab = topology.build_halo_table(N, p, g) if table is None else table
    n = N // p
    m = n + 2 * g
    pp = p * p
    delta = 0.5 * np.pi / N
    text = extended_tangents(N, g)
    dst_face, src_face = tab["dst_tile"] // pp, tab["src_tile"] // pp
    rd, rs = tab["dst_tile"] % pp, tab["src_tile"] % pp
    # face-level indices (interior numbering, ghosts negative or >= N) of the ghost cell and of the copy's source
    J = (rd // p) * n + tab["dst_j"] - g
    I = (rd % p) * n + tab["dst_i"] - g
    Js = (rs // p) * n + tab["src_j"] - g
    Is = (rs % p) * n + tab["src_i"] - g
    count = J.size
    dst = (tab["dst_tile"] * m + tab["dst_j"]) * m + tab["dst_i"]
    copy_src = (tab["src_tile"] * m + tab["src_j"]) * m + tab["src_i"]
    src = np.repeat(copy_src[:, None], 4, axis=1)
    w = np.zeros((count, 4))
    w[:, 0] = 1.0
    code = tab["code"]
    rot = np.zeros((count, 2, 2))
    swap = (code & 1) == 1
    rot[~swap, 0, 0] = rot[~swap, 1, 1] = 1.0
    rot[swap, 0, 1] = rot[swap, 1, 0] = 1.0
    rot[(code & 2) != 0, 0, :] *= -1.0
    rot[(code & 4) != 0, 1, :] *= -1.0
    # seam edge strips: exactly one face-level index outside [0, N), and the source on another face
    out_j, out_i = (J < 0) | (J >= N), (I < 0) | (I >= N)
    sel = np.nonzero((out_j ^ out_i) & (dst_face != src_face))[0]
    if sel.size:
        L, B = dst_face[sel], src_face[sel]
        fr = np.asarray(FACE_FRAMES, dtype=np.float64)
        X, Y = text[I[sel] + g], text[J[sel] + g]
        P = fr[L, 0] + X[:, None] * fr[L, 1] + Y[:, None] * fr[L, 2]
        P /= np.linalg.norm(P, axis=-1, keepdims=True)
        _, _, XB, YB = chart_basis(B, P, a)
        fi = (np.arctan(XB) + 0.25 * np.pi) / delta - 0.5          # fractional interior indices in the neighbour
        fj = (np.arctan(YB) + 0.25 * np.pi) / delta - 0.5
        # the copy's source sits at the same depth: that index is the integer one, the other is interpolated along
        along_j = np.abs(fi - Is[sel]) <= np.abs(fj - Js[sel])
        depth_err = np.where(along_j, np.abs(fi - Is[sel]), np.abs(fj - Js[sel]))
        if depth_err.max() > 1e-6:
            raise RuntimeError("a ghost point does not lie on its neighbour's coordinate line (not an equiangular grid?)")
        f = np.where(along_j, fj, fi)
        b0 = np.clip(np.floor(f).astype(np.int64) - (points // 2 - 1), 0, N - points)
        nodes = b0[:, None] + np.arange(4)[None, :]
        nodes = np.minimum(nodes, (b0 + points - 1)[:, None])      # unused slots repeat the last node with weight 0
        ws = np.zeros((sel.size, 4))
        for q in range(points):
            lag = np.ones(sel.size)
            for r in range(points):
                if r != q:
                    lag *= (f - nodes[:, r]) / (nodes[:, q] - nodes[:, r])
            ws[:, q] = lag
        jb = np.where(along_j[:, None], nodes, Js[sel][:, None])
        ib = np.where(along_j[:, None], Is[sel][:, None], nodes)
        tile = B[:, None] * pp + (jb // n) * p + (ib // n)
        src[sel] = (tile * m + jb % n + g) * m + ib % n + g
        w[sel] = ws
        rot[sel] = seam_rotation(L, B, XB, YB, "contravariant", a)
    return {"dst": dst, "src": src, "w": w, "rot": rot, "tangents": text}
