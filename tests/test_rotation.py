"""Row a9: host-precomputed geometric seam-rotation tables for the vector halo (geometry.seam_rotation_tables).

The reference exchanges Cartesian velocity components to side-step this (deck slide 18, "Cartesian Velocity Exchange");
BASELINE.json keeps panel-local components and rotates them at the seams.  CPU-only checks of the generator; the GPU
check with the same tables is in test_gpu_round2.py.
"""
import numpy as np
import pytest

from cubesphere import geometry as geo
from cubesphere import topology


def _solid_body(P, omega=(0.3, -0.5, 0.8), a=geo.EARTH_RADIUS):
    """Cartesian velocity of a rigid rotation about a tilted axis at unit points P [..., 3] (m/s)."""
    w = np.asarray(omega, dtype=np.float64) * 1e-5
    return a * np.cross(np.broadcast_to(w, P.shape), P)


def _contra(face, P, V):
    ea, eb, X, Y = geo.chart_basis(face, P)
    g11, g12, g22 = np.sum(ea * ea, -1), np.sum(ea * eb, -1), np.sum(eb * eb, -1)
    b1, b2 = np.sum(ea * V, -1), np.sum(eb * V, -1)
    det = g11 * g22 - g12 * g12
    return (g22 * b1 - g12 * b2) / det, (g11 * b2 - g12 * b1) / det, b1, b2


def _table_points(tab, N, p, g):
    n = N // p
    t = geo.centre_tangents(N)
    pp = p * p
    sf, df = tab["src_tile"] // pp, tab["dst_tile"] // pp
    r = tab["src_tile"] % pp
    X = t[(r % p) * n + tab["src_i"] - g]
    Y = t[(r // p) * n + tab["src_j"] - g]
    fr = np.asarray(topology.FACE_FRAMES, dtype=np.float64)[sf]
    P = fr[:, 0] + X[:, None] * fr[:, 1] + Y[:, None] * fr[:, 2]
    return sf, df, P / np.linalg.norm(P, axis=-1, keepdims=True)


@pytest.mark.parametrize("N,p,g", [(24, 1, 2), (16, 2, 2), (12, 1, 3)])
@pytest.mark.parametrize("kind", ["contravariant", "covariant"])
def test_rotated_ghosts_equal_the_analytic_field_in_the_local_extended_chart(N, p, g, kind):
    """A solid-body wind, held by every panel in its own chart: after R, the neighbour's components are the
    field's components in OUR extended chart at the neighbour's cell centre, to round-off."""
    tab = topology.build_halo_table(N, p, g)
    R = geo.seam_rotation_tables(N, g, p, kind)
    assert R.shape == (tab["dst_tile"].size, 2, 2)
    sf, df, P = _table_points(tab, N, p, g)
    V = _solid_body(P)
    us, vs, c1s, c2s = _contra(sf, P, V)                  # what the neighbour holds
    ul, vl, c1l, c2l = _contra(df, P, V)                  # what we want in the ghost cell
    src = np.stack([us, vs], -1) if kind == "contravariant" else np.stack([c1s, c2s], -1)
    want = np.stack([ul, vl], -1) if kind == "contravariant" else np.stack([c1l, c2l], -1)
    got = np.einsum("kij,kj->ki", R, src)
    scale = np.abs(want).max()
    assert np.abs(got - want).max() <= 1e-12 * scale
    # inside a face (tile-to-tile ghosts) nothing is rotated: the exact identity
    same = sf == df
    if p > 1:
        assert same.any()
    assert np.array_equal(R[same], np.broadcast_to(np.eye(2), R[same].shape))


def test_rotation_tends_to_the_signed_permutation_at_edge_centres():
    """Near the middle of a seam the two charts' coordinate lines line up, so R approaches the seam's signed
    permutation (what the default copy halo applies everywhere); towards the cube corners it departs from it by O(1)."""
    N, g = 64, 2
    tab = topology.build_halo_table(N, 1, g)
    R = geo.seam_rotation_tables(N, g)
    code = tab["code"]
    S = np.zeros_like(R)
    swap = (code & 1) == 1
    S[~swap, 0, 0] = 1.0
    S[~swap, 1, 1] = 1.0
    S[swap, 0, 1] = 1.0
    S[swap, 1, 0] = 1.0
    S[(code & 2) != 0, 0, :] *= -1.0
    S[(code & 4) != 0, 1, :] *= -1.0
    dev = np.abs(R - S).max(axis=(1, 2))
    # along-edge position of the source cell: distance from the edge centre in cells
    along = np.where((tab["dst_j"] < g) | (tab["dst_j"] >= g + N), tab["dst_i"], tab["dst_j"]) - g
    centre = np.abs(along - (N - 1) / 2) < 1.0
    edge = (along >= 0) & (along < N)
    first = edge & centre & (np.minimum(np.minimum(tab["dst_j"], tab["dst_i"]), 2 * g + N - 1 - np.maximum(tab["dst_j"], tab["dst_i"])) == g - 1)
    assert first.any() and dev[first].max() < 0.06        # O(delta) at depth 0 next to the edge centre
    assert dev[edge].max() > 0.3                           # O(1) next to the cube corners


def test_rotation_table_argument_checks():
    with pytest.raises(ValueError):
        geo.seam_rotation_tables(8, 2, kind="cartesian")
    with pytest.raises(ValueError):                        # the antipodal face is not in the chart's hemisphere
        geo.chart_basis(0, np.array([[0.0, 0.0, -1.0]]))
