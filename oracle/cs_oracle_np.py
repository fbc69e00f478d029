"""TEST INFRASTRUCTURE -- CPU oracle (NumPy fp64) for the cubed-sphere shallow-water step.

PARITY UNPINNED UPSTREAM: the reference (areanddee/Sharding-the-Sphere-Fall-2025-
JAX-DevLab-Examples) ships no tests, golden vectors, fixtures or runnable solver,
and JAX is not installed here, so nothing of the reference can be executed to pin
this oracle.  The only reference data that can be checked are the 12 seam
pairings with their op flags (JAX-DevLab-Examples.py:114-139); tests/ checks
:func:`adjacency` against that table.  Everything else is self-validated
(panel areas sum to 4 pi a^2, seam-flux antisymmetry, mass to round-off, TC2
steadiness, TC1 bell transport) -- see tests/test_oracle.py.

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline / reference arm
may import this module.  The product path never does.

What it restates, and from where:
  * layout ``field[face, j, i]`` with a ghost ring, interior ``[g:g+N, g:g+N]``,
    ghosts filled from the *neighbour's interior* cells
    (reference JAX-DevLab-Examples.py:141,174,183-195);
  * 'R' = reverse the along-edge index, 'T' = the two edges run along different
    array axes (JAX-DevLab-Examples.py:143-163); the same op is applied in both
    directions of a seam (JAX-DevLab-Examples.py:188-189);
  * everything numerical (PLR upwind continuity, vector-invariant momentum,
    SSP-RK, diagnostics) is a clean-room scheme from the literature, specified
    in DESIGN.md section 3; the reference has no numerics.

The halo fill here is written with whole-strip slicing per seam, on purpose
unlike the product's per-ghost-cell gather table, so the two check each other.
"""
from __future__ import annotations

import numpy as np

EARTH_RADIUS = 6.37122e6
EARTH_OMEGA = 7.292e-5
EARTH_GRAVITY = 9.80616

# (normal, local x, local y): unique right-handed frames reproducing the reference's flags
FRAMES = np.array([
    [[0, 0, 1], [1, 0, 0], [0, 1, 0]],
    [[0, 1, 0], [-1, 0, 0], [0, 0, 1]],
    [[-1, 0, 0], [0, -1, 0], [0, 0, 1]],
    [[0, -1, 0], [1, 0, 0], [0, 0, 1]],
    [[1, 0, 0], [0, 1, 0], [0, 0, 1]],
    [[0, 0, -1], [-1, 0, 0], [0, 1, 0]],
], dtype=np.int64)

# Shu-Osher pairs (a, b): q_next = a q0 + b (q_s + dt L(q_s)).  a is stored as 1 - b so that
# a + b == 1 exactly in binary floating point (1/3 + 2/3 rounds to 1 - 2^-54 and leaks mass).
RK_SCHEMES = {
    "euler": ((0.0, 1.0),),
    "ssprk2": ((0.0, 1.0), (0.5, 0.5)),
    "ssprk3": ((0.0, 1.0), (0.75, 0.25), (1.0 - 2.0 / 3.0, 2.0 / 3.0)),
}
LIMITERS = ("none", "minmod", "mc")


# ----------------------------------------------------------------------------
# topology
# ----------------------------------------------------------------------------

def _outward(f, e):
    _, x, y = FRAMES[f]
    return {"N": y, "S": -y, "E": x, "W": -x}[e]


def _along(f, e):
    _, x, y = FRAMES[f]
    return x if e in "NS" else y


def adjacency():
    """{(face, edge): (face_b, edge_b, reverse, S)} from the frames alone."""
    adj = {}
    for f in range(6):
        for e in "NSEW":
            o = _outward(f, e)
            fb = next(k for k in range(6) if (FRAMES[k, 0] == o).all())
            eb = next(k for k in "NSEW" if (_outward(fb, k) == FRAMES[f, 0]).all())
            t = _along(f, e)
            rev = int(t @ _along(fb, eb)) < 0
            S = np.zeros((2, 2), dtype=np.int64)
            for col in (0, 1):
                w = FRAMES[fb, 1 + col]
                img = int(w @ t) * t - int(w @ _outward(fb, eb)) * o
                S[:, col] = (int(img @ FRAMES[f, 1]), int(img @ FRAMES[f, 2]))
            adj[(f, e)] = (fb, eb, rev, S)
    return adj


_ADJ = adjacency()


def op_flag(f, e):
    fb, eb, rev, _ = _ADJ[(f, e)]
    return (("T" if (e in "NS") != (eb in "NS") else "") + ("R" if rev else "")) or "N"


# ----------------------------------------------------------------------------
# halo fill
# ----------------------------------------------------------------------------

def _strip(a, e, N, g):
    """Interior strip of one face next to edge e in canonical form S[d, k]."""
    if e == "N":
        return a[g + N - g:g + N, g:g + N][::-1, :]
    if e == "S":
        return a[g:2 * g, g:g + N]
    if e == "E":
        return a[g:g + N, N:g + N].T[::-1, :]
    return a[g:g + N, g:2 * g].T


def _set_ghost(a, e, G, N, g):
    """Write canonical strip G[d, k] into the ghost cells of edge e."""
    if e == "N":
        a[g + N:g + N + g, g:g + N] = G
    elif e == "S":
        a[0:g, g:g + N] = G[::-1, :]
    elif e == "E":
        a[g:g + N, g + N:g + N + g] = G.T
    else:
        a[g:g + N, 0:g] = G[::-1, :].T


def _fill_corners(a, N, g):
    a[:, 0:g, 0:g] = a[:, g:g + 1, 0:g]
    a[:, 0:g, g + N:] = a[:, g:g + 1, g + N:]
    a[:, g + N:, 0:g] = a[:, g + N - 1:g + N, 0:g]
    a[:, g + N:, g + N:] = a[:, g + N - 1:g + N, g + N:]


def halo_fill_scalar(field, g):
    """Fill all ghost cells of ``field[6, N+2g, N+2g]``; returns a new array."""
    out = np.array(field, copy=True)
    N = field.shape[-1] - 2 * g
    for (f, e), (fb, eb, rev, _) in _ADJ.items():
        G = _strip(field[fb], eb, N, g)
        if rev:
            G = G[:, ::-1]
        _set_ghost(out[f], e, G, N, g)
    _fill_corners(out, N, g)
    return out


def halo_fill_vector(u, v, g):
    """Ghost fill of panel-local vector components: copy + index reversal + the
    seam's signed permutation S of the two components.  Returns new (u, v)."""
    ou, ov = np.array(u, copy=True), np.array(v, copy=True)
    N = u.shape[-1] - 2 * g
    for (f, e), (fb, eb, rev, S) in _ADJ.items():
        Gu, Gv = _strip(u[fb], eb, N, g), _strip(v[fb], eb, N, g)
        if rev:
            Gu, Gv = Gu[:, ::-1], Gv[:, ::-1]
        src = (Gu, Gv)
        comps = []
        for row in (0, 1):
            col = 0 if S[row, 0] != 0 else 1
            comps.append(-src[col] if S[row, col] < 0 else src[col])
        _set_ghost(ou[f], e, comps[0], N, g)
        _set_ghost(ov[f], e, comps[1], N, g)
    _fill_corners(ou, N, g)
    _fill_corners(ov, N, g)
    return ou, ov


def halo_fill_state(h, u, v, g):
    h2 = halo_fill_scalar(h, g)
    u2, v2 = halo_fill_vector(u, v, g)
    return h2, u2, v2


# ----------------------------------------------------------------------------
# geometry tables
# ----------------------------------------------------------------------------

def centre_tangents(N):
    i = np.arange(N, dtype=np.float64)
    x = np.tan(-0.25 * np.pi + (i + 0.5) * (0.5 * np.pi / N))
    half = N // 2
    x[:half] = -x[::-1][:half]
    if N % 2:
        x[half] = 0.0
    return x


def padded_tangents(N, g):
    return centre_tangents(N)[np.arange(-g, N + g) % N]


class Params:
    """Scheme parameters shared by every oracle entry point."""

    def __init__(self, N, g=2, a=EARTH_RADIUS, omega=EARTH_OMEGA, grav=EARTH_GRAVITY,
                 axis=(0.0, 0.0, 1.0), limiter="mc", rk="ssprk3", advect_only=False, hs=None, tangents=None, halo_fill=None):
        """``tangents``: optional table of the N + 2g tangents tan(alpha_I) incl. ghosts (default: the periodic
        continuation); ``halo_fill``: optional replacement ``f(h, u, v) -> (h, u, v)`` for the ghost
        fill after each stage.  Both are hooks for oracle/seam_study.py; the parity oracle uses neither."""
        if g < 2:
            raise ValueError("the step needs a halo of width g >= 2")
        if limiter not in LIMITERS:
            raise ValueError(f"Unknown limiter: {limiter}")
        if rk not in RK_SCHEMES:
            raise ValueError(f"Unknown rk scheme: {rk}")
        self.N, self.g, self.a, self.omega, self.grav = N, g, a, omega, grav
        self.axis = np.asarray(axis, dtype=np.float64)
        self.limiter, self.rk, self.advect_only = limiter, rk, advect_only
        self.delta = 0.5 * np.pi / N
        t = padded_tangents(N, g) if tangents is None else np.asarray(tangents, dtype=np.float64)
        self.halo_fill = halo_fill
        X = t[None, :]
        Y = t[:, None]
        A = 1.0 + X * X
        B = 1.0 + Y * Y
        rho2 = A + Y * Y
        r2inv = 1.0 / rho2
        c = (a * a) * A * B * (r2inv * r2inv)
        self.sg = c * np.sqrt(rho2)
        self.g11, self.g12, self.g22 = c * A, -(c * X) * Y, c * B
        ia2 = 1.0 / (a * a)
        self.gi11 = (rho2 * ia2) / A
        self.gi12 = ((rho2 * ia2) * (X * Y)) / (A * B)
        self.gi22 = (rho2 * ia2) / B
        cf = np.array([[self.axis @ FRAMES[f, k] for k in range(3)] for f in range(6)], dtype=np.float64)
        # f * sqrt(G) = 2 Omega (c0 + c1 X + c2 Y) / rho * sqrt(G) = 2 Omega (c0 + c1 X + c2 Y) c
        self.fsg = (2.0 * omega) * (cf[:, 0, None, None] + cf[:, 1, None, None] * X + cf[:, 2, None, None] * Y) * c
        M = N + 2 * g
        if hs is None:
            self.hs = np.zeros((6, M, M))
        else:
            hs = np.asarray(hs, dtype=np.float64)
            if hs.shape == (6, N, N):
                full = np.zeros((6, M, M))
                full[:, g:g + N, g:g + N] = hs
                hs = halo_fill_scalar(full, g)
            self.hs = hs


def _limited_slope(dm, dp, limiter):
    if limiter == "none":
        return 0.5 * (dm + dp)
    sgn = 0.5 * (np.sign(dm) + np.sign(dp))
    if limiter == "minmod":
        return sgn * np.minimum(np.abs(dm), np.abs(dp))
    return sgn * np.minimum(np.minimum(2.0 * np.abs(dm), 2.0 * np.abs(dp)), 0.5 * np.abs(dm + dp))


def tendency(p: Params, h, u, v):
    """L(q) on interior cells from halo-filled ``h, u, v [6, M, M]``: returns (dh, du, dv) [6, N, N]."""
    N, g = p.N, p.g
    M = N + 2 * g
    m1 = p.sg * u
    m2 = p.sg * v
    c1 = p.g11 * u + p.g12 * v
    c2 = p.g12 * u + p.g22 * v
    K = 0.5 * (u * c1 + v * c2)
    Bn = p.grav * (h + p.hs) + K
    rd = 1.0 / p.delta
    r2d = 0.5 / p.delta
    c = slice(g, g + N)

    # x direction: slopes on columns 1..M-2, faces between I and I+1 for I = 1..M-3
    sx = np.zeros_like(h)
    sx[:, :, 1:M - 1] = _limited_slope(h[:, :, 1:M - 1] - h[:, :, 0:M - 2], h[:, :, 2:M] - h[:, :, 1:M - 1], p.limiter)
    hL = h[:, :, 1:M - 2] + 0.5 * sx[:, :, 1:M - 2]
    hR = h[:, :, 2:M - 1] - 0.5 * sx[:, :, 2:M - 1]
    Fm = 0.5 * (m1[:, :, 1:M - 2] + m1[:, :, 2:M - 1])
    F = np.maximum(Fm, 0.0) * hL + np.minimum(Fm, 0.0) * hR          # F[..., k] is the face right of column k+1
    # interior column I has right face index I-1 and left face index I-2
    dFx = F[:, c, g - 1:g + N - 1] - F[:, c, g - 2:g + N - 2]

    sy = np.zeros_like(h)
    sy[:, 1:M - 1, :] = _limited_slope(h[:, 1:M - 1, :] - h[:, 0:M - 2, :], h[:, 2:M, :] - h[:, 1:M - 1, :], p.limiter)
    hS = h[:, 1:M - 2, :] + 0.5 * sy[:, 1:M - 2, :]
    hN = h[:, 2:M - 1, :] - 0.5 * sy[:, 2:M - 1, :]
    Gm = 0.5 * (m2[:, 1:M - 2, :] + m2[:, 2:M - 1, :])
    G = np.maximum(Gm, 0.0) * hS + np.minimum(Gm, 0.0) * hN
    dGy = G[:, g - 1:g + N - 1, c] - G[:, g - 2:g + N - 2, c]

    dh = -((dFx + dGy) * rd) * (1.0 / p.sg[c, c])
    if p.advect_only:
        z = np.zeros_like(dh)
        return dh, z, z

    eta = ((c2[:, c, g + 1:g + N + 1] - c2[:, c, g - 1:g + N - 1])
           - (c1[:, g + 1:g + N + 1, c] - c1[:, g - 1:g + N - 1, c])) * r2d + p.fsg[:, c, c]
    T1 = eta * v[:, c, c] - (Bn[:, c, g + 1:g + N + 1] - Bn[:, c, g - 1:g + N - 1]) * r2d
    T2 = -eta * u[:, c, c] - (Bn[:, g + 1:g + N + 1, c] - Bn[:, g - 1:g + N - 1, c]) * r2d
    du = p.gi11[c, c] * T1 + p.gi12[c, c] * T2
    dv = p.gi12[c, c] * T1 + p.gi22[c, c] * T2
    return dh, du, dv


def rk_stage(p: Params, q0, qs, a, b, dt):
    """out = a q0 + b (qs + dt L(qs)) on interior cells, then halo fill.  q = (h, u, v), halo-filled."""
    g, N = p.g, p.N
    c = slice(g, g + N)
    d = tendency(p, *qs)
    out = []
    for x0, xs, dx in zip(q0, qs, d):
        o = np.zeros_like(xs)
        o[:, c, c] = a * x0[:, c, c] + b * (xs[:, c, c] + dt * dx)
        out.append(o)
    if p.halo_fill is not None:
        return p.halo_fill(out[0], out[1], out[2])
    return halo_fill_state(out[0], out[1], out[2], g)


def step(p: Params, state, dt):
    """One explicit time step; ``state = (h, u, v)`` halo-filled ``[6, M, M]``; returns the new state."""
    q0 = state
    qs = state
    for a, b in RK_SCHEMES[p.rk]:
        qs = rk_stage(p, q0, qs, a, b, dt)
    return qs


def diagnostics(p: Params, h, u, v):
    """(mass, total energy): sums over interior cells of sqrt(G) {h, h K + g h^2/2 + g h hs} dalpha dbeta."""
    g, N = p.g, p.N
    c = slice(g, g + N)
    hh, uu, vv = h[:, c, c], u[:, c, c], v[:, c, c]
    c1 = p.g11[c, c] * uu + p.g12[c, c] * vv
    c2 = p.g12[c, c] * uu + p.g22[c, c] * vv
    K = 0.5 * (uu * c1 + vv * c2)
    w = p.sg[c, c] * (p.delta * p.delta)
    mass = float(np.sum(w * hh))
    energy = float(np.sum(w * (hh * K + 0.5 * p.grav * hh * hh + p.grav * hh * p.hs[:, c, c])))
    return mass, energy


# ----------------------------------------------------------------------------
# helpers for tests
# ----------------------------------------------------------------------------

def pad(field, g):
    """[6, N, N] -> [6, N+2g, N+2g] with zero ghosts."""
    N = field.shape[-1]
    out = np.zeros((field.shape[0], N + 2 * g, N + 2 * g), dtype=field.dtype)
    out[:, g:g + N, g:g + N] = field
    return out


def faces_to_tiles(field, p, g):
    """Halo-filled faces [6, N+2g, N+2g] -> tiles [6 p^2, n+2g, n+2g] (ghosts taken from the face array)."""
    N = field.shape[-1] - 2 * g
    n = N // p
    out = np.empty((6 * p * p, n + 2 * g, n + 2 * g), dtype=field.dtype)
    for f in range(6):
        for ty in range(p):
            for tx in range(p):
                out[f * p * p + ty * p + tx] = field[f, ty * n:ty * n + n + 2 * g, tx * n:tx * n + n + 2 * g]
    return out


def tiles_to_faces(tiles, p, g):
    """Tiles [6 p^2, n+2g, n+2g] -> faces [6, N+2g, N+2g]; only interiors are assembled (ghosts zero)."""
    n = tiles.shape[-1] - 2 * g
    N = n * p
    out = np.zeros((6, N + 2 * g, N + 2 * g), dtype=tiles.dtype)
    for f in range(6):
        for ty in range(p):
            for tx in range(p):
                out[f, g + ty * n:g + ty * n + n, g + tx * n:g + tx * n + n] = \
                    tiles[f * p * p + ty * p + tx, g:g + n, g:g + n]
    return out


# ----------------------------------------------------------------------------
# demo workload: scalar diffusion (the reference's heat-diffusion "proof that sharding works", deck slides 12, 17)
# ----------------------------------------------------------------------------

def diffuse(p: Params, T, kappa_dt, source=None, dt_source=0.0):
    """One explicit step of the 5-point flux-form diffusion of a ghost-filled scalar ``T [6, M, M]``:
    T + kappa_dt / (sqrt(G) delta^2) [d_a(W1 d_a T) + d_b(W2 d_b T)] + dt_source * source, W1 = sqrt(G) g^11 and
    W2 = sqrt(G) g^22 averaged to the faces (no cross-derivative terms; see csrc/cs_scalar.cu).  Ghosts refilled."""
    g, N = p.g, p.N
    c = slice(g, g + N)
    w1, w2 = p.sg * p.gi11, p.sg * p.gi22
    FE = 0.5 * (w1[c, g:g + N] + w1[c, g + 1:g + N + 1]) * (T[:, c, g + 1:g + N + 1] - T[:, c, c])
    FW = 0.5 * (w1[c, g - 1:g + N - 1] + w1[c, g:g + N]) * (T[:, c, c] - T[:, c, g - 1:g + N - 1])
    GN = 0.5 * (w2[g:g + N, c] + w2[g + 1:g + N + 1, c]) * (T[:, g + 1:g + N + 1, c] - T[:, c, c])
    GS = 0.5 * (w2[g - 1:g + N - 1, c] + w2[g:g + N, c]) * (T[:, c, c] - T[:, g - 1:g + N - 1, c])
    out = np.zeros_like(T)
    out[:, c, c] = T[:, c, c] + (kappa_dt / (p.delta * p.delta)) * (((FE - FW) + (GN - GS)) / p.sg[c, c])
    if source is not None:
        out[:, c, c] += dt_source * source[:, c, c]
    return halo_fill_scalar(out, g)


def checkerboard_source(N, g, squares=4, amplitude=1.0):
    """The demo's heat source: a checkerboard of ``squares`` x ``squares`` patches on the top panel (face 0)."""
    src = np.zeros((6, N + 2 * g, N + 2 * g))
    k = (np.arange(N) * squares) // N
    src[0, g:g + N, g:g + N] = amplitude * (((k[:, None] + k[None, :]) % 2) == 0)
    return src
