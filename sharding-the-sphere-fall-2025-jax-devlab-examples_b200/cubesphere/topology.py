"""Cube topology: faces, the 12 seams, tiles and the ghost-cell resolver.

Facts taken from the reference (conventions only, no code):
  * face numbering 0..5, edge names N/S/E/W, the 12 pairings with their
    ``N``/``T``/``R``/``TR`` flags and their grouping into 4 stages
    (reference JAX-DevLab-Examples.py:105-139, slide 9 of the deck);
  * ghost <- neighbour *interior* semantics (JAX-DevLab-Examples.py:183-195);
  * layout ``(tiles, y + ghost, x + ghost)`` (JAX-DevLab-Examples.py:141,174).

Everything else (frames, the resolver, tiles with tiles_per_edge > 1, the
signed-permutation of vector components, the cube-corner rule) is derived
here from plain cube geometry.  The frames below are the unique right-handed
assignment that reproduces every flag of the reference table (checked in
tests/test_topology.py against the table itself).

Array convention: ``field[tile, j, i]`` -- axis 1 is local y (S -> N), axis 2 is
local x (W -> E, fastest).  Interior cells are ``g <= i, j < g + n``.
"""
from __future__ import annotations

import numpy as np

# (outward normal, local x, local y) as integer 3-vectors; x cross y == normal.
FACE_FRAMES = (
    ((0, 0, 1), (1, 0, 0), (0, 1, 0)),    # 0: +Z  north-polar panel
    ((0, 1, 0), (-1, 0, 0), (0, 0, 1)),   # 1: +Y  lon  90
    ((-1, 0, 0), (0, -1, 0), (0, 0, 1)),  # 2: -X  lon 180
    ((0, -1, 0), (1, 0, 0), (0, 0, 1)),   # 3: -Y  lon 270
    ((1, 0, 0), (0, 1, 0), (0, 0, 1)),    # 4: +X  lon   0
    ((0, 0, -1), (-1, 0, 0), (0, 1, 0)),  # 5: -Z  south-polar panel
)
EDGES = ("N", "S", "E", "W")
EDGE_ID = {e: k for k, e in enumerate(EDGES)}

# Stage in which each seam is exchanged by the reference's 4-stage schedule
# (JAX-DevLab-Examples.py:114-139).  Only the *staging* is recorded here; the
# partner edge and the op flag of every entry are derived from FACE_FRAMES.
_STAGE_ORDER = (
    ((0, "N"), (3, "E"), (2, "S")),
    ((0, "E"), (2, "E"), (1, "S")),
    ((0, "W"), (1, "W"), (3, "S")),
    ((0, "S"), (1, "E"), (4, "S")),
)


def _vec(v):
    return np.asarray(v, dtype=np.int64)


def edge_outward(face: int, edge: str) -> np.ndarray:
    """3-D unit vector pointing out of ``face`` across ``edge`` (in the face plane)."""
    _, x, y = FACE_FRAMES[face]
    return {"N": _vec(y), "S": -_vec(y), "E": _vec(x), "W": -_vec(x)}[edge]


def edge_along(face: int, edge: str) -> np.ndarray:
    """3-D direction of increasing along-edge index (x for N/S edges, y for E/W)."""
    _, x, y = FACE_FRAMES[face]
    return _vec(x) if edge in ("N", "S") else _vec(y)


def neighbour(face: int, edge: str):
    """Return ``(face_b, edge_b, reverse, S)`` for the seam at (face, edge).

    ``reverse`` -- the along-edge indices of the two faces run anti-parallel
    (the reference's ``R`` flag, JAX-DevLab-Examples.py:159-162).
    ``S`` -- 2x2 signed permutation: components of a tangent vector expressed
    in the neighbour's local (x, y) axes, carried across the seam by unfolding
    the two faces flat, become ``S @ (u, v)`` in this face's local axes.
    """
    n_a, x_a, y_a = (_vec(v) for v in FACE_FRAMES[face])
    o_a = edge_outward(face, edge)
    t_a = edge_along(face, edge)
    for fb in range(6):
        if np.array_equal(_vec(FACE_FRAMES[fb][0]), o_a):
            break
    else:  # pragma: no cover
        raise AssertionError("no neighbour found")
    for eb in EDGES:
        if np.array_equal(edge_outward(fb, eb), n_a):
            break
    t_b = edge_along(fb, eb)
    reverse = int(t_a @ t_b) < 0
    o_b = edge_outward(fb, eb)          # == n_a
    _, x_b, y_b = (_vec(v) for v in FACE_FRAMES[fb])
    S = np.zeros((2, 2), dtype=np.int64)
    for col, w in enumerate((x_b, y_b)):
        image = int(w @ t_a) * t_a + int(w @ o_b) * (-o_a)   # unfold into A's plane
        S[0, col] = int(image @ x_a)
        S[1, col] = int(image @ y_a)
    return fb, eb, reverse, S


def seam_op(face: int, edge: str) -> str:
    """The reference's op string for the seam: 'N', 'T', 'R' or 'TR'."""
    fb, eb, reverse, _ = neighbour(face, edge)
    transpose = (edge in ("N", "S")) != (eb in ("N", "S"))
    return (("T" if transpose else "") + ("R" if reverse else "")) or "N"


def create_communication_schedule():
    """12 seam swaps in 4 conflict-free stages.

    Same return shape as the reference function of the same name
    (JAX-DevLab-Examples.py:105-139): a tuple of 4 stages, each a tuple of 3
    ``((face_a, edge_a), (face_b, edge_b), ops)`` entries; every stage is a
    perfect matching of the six faces.
    """
    schedule = []
    for stage in _STAGE_ORDER:
        entries = []
        for face, edge in stage:
            fb, eb, _, _ = neighbour(face, edge)
            entries.append(((face, edge), (fb, eb), seam_op(face, edge)))
        schedule.append(tuple(entries))
    return tuple(schedule)


def create_tile_schedule(tiles_per_edge: int = 1):
    """Edge-coloured exchange schedule for 6 p^2 tiles: the "scalable edge coloring" generalisation the
    reference names but does not ship (deck slide 9; JAX-DevLab-Examples.py:31-37 raises for p != 1).

    Returns a tuple of stages; each stage is a tuple of ``((tile_a, edge_a), (tile_b, edge_b), ops)``
    entries in which no tile appears twice, and every one of the 12 p^2 tile-edge pairs (seams and
    interior edges) appears exactly once.  ``ops`` is 'N' inside a face and the seam flag otherwise.
    For p = 1 the result is :func:`create_communication_schedule` (the reference's own 4 stages).
    The tile graph is 4-regular; colouring is greedy over a fixed edge order with repair by Kempe-chain
    swaps, so it uses 4 stages whenever the greedy pass does and never more than 5 (Vizing).
    The B200 halo kernels do not need the schedule (each ghost cell has a single writer); it exists for
    callers that exchange by pairwise swaps like the reference does.
    """
    p = int(tiles_per_edge)
    if p < 1:
        raise ValueError(f"tiles_per_edge must be a positive integer, got {p}")
    if p == 1:
        return create_communication_schedule()
    desc = edge_push_descriptors(p)
    pairs = []
    seen = set()
    for t in range(num_tiles(p)):
        face, ty, tx = tile_coords(t, p)
        for e, name in enumerate(EDGES):
            dt_, de = int(desc[t, e, 0]), int(desc[t, e, 1])
            key = frozenset(((t, e), (dt_, de)))
            if key in seen:
                continue
            seen.add(key)
            inside = {"N": ty < p - 1, "S": ty > 0, "E": tx < p - 1, "W": tx > 0}[name]
            pairs.append(((t, name), (dt_, EDGES[de]), "N" if inside else seam_op(face, name)))
    colour = {}
    used = {t: {} for t in range(num_tiles(p))}          # tile -> {colour: pair index}
    ncol = 4

    def free(t):
        return [c for c in range(ncol) if c not in used[t]]

    for k, ((ta, _), (tb, _), _) in enumerate(pairs):
        while True:
            common = [c for c in free(ta) if c in free(tb)]
            if common:
                c = common[0]
                break
            fa, fb = free(ta), free(tb)
            if not fa or not fb:
                ncol += 1
                continue
            # Kempe chain from tb alternating colours (ca, cb): swap them so that ca becomes free at tb
            ca, cb = fa[0], fb[0]
            chain, node, want = [], tb, ca
            while want in used[node]:
                j = used[node][want]
                chain.append(j)
                (x, _), (y, _), _ = pairs[j]
                node = y if x == node else x
                want = cb if want == ca else ca
            if node == ta:                                 # the chain ends at ta: swapping would not help
                ncol += 1
                continue
            for j in chain:
                (x, _), (y, _), _ = pairs[j]
                old = colour[j]
                del used[x][old], used[y][old]
            for j in chain:
                (x, _), (y, _), _ = pairs[j]
                new = cb if colour[j] == ca else ca
                colour[j] = new
                used[x][new] = j
                used[y][new] = j
        colour[k] = c
        used[ta][c] = k
        used[tb][c] = k
    stages = [[] for _ in range(ncol)]
    for k, pr in enumerate(pairs):
        stages[colour[k]].append(pr)
    return tuple(tuple(st) for st in stages if st)


def apply_operations(data, operations: str):
    """Reverse the along-edge axis (last axis) for 'R'/'TR'; identity for 'N'/'T'.

    Mirrors reference JAX-DevLab-Examples.py:143-163, including the ValueError
    for an unknown flag.  ``T`` is absorbed by extracting strips in
    edge-canonical (depth, along-edge) form, as in the reference where the data
    are 1-D.
    """
    if operations in ("N", "T"):
        return data
    if operations in ("R", "TR"):
        return data.flip(-1) if hasattr(data, "flip") else data[..., ::-1]
    raise ValueError(f"seam operation must be one of N, T, R, TR; got {operations!r}")


def perm_code(S: np.ndarray) -> int:
    """Encode a 2x2 signed permutation: bit0 swap, bit1 negate out-u, bit2 negate out-v."""
    swap = int(S[0, 0] == 0)
    s1 = S[0, 1] if swap else S[0, 0]
    s2 = S[1, 0] if swap else S[1, 1]
    assert abs(s1) == 1 and abs(s2) == 1
    return swap | (int(s1 < 0) << 1) | (int(s2 < 0) << 2)


def resolve(face: int, J: np.ndarray, I: np.ndarray, N: int):
    """Where does the value of face-level cell (J, I) of ``face`` come from?

    ``J, I`` are integer arrays of face-level interior coordinates in
    ``[-g, N + g)`` (negative / >= N means ghost).  Returns ``(fb, Jb, Ib, code)``
    arrays: source face, source interior coordinates in ``[0, N)`` and the
    vector-component code (see :func:`perm_code`).

    Cells off both axes (the g x g blocks at the 8 cube vertices) have no
    geometric owner; the rule is constant extrapolation of the E/W ghost column
    in y, i.e. J is clamped into [0, N) first.
    """
    J = np.asarray(J, dtype=np.int64)
    I = np.asarray(I, dtype=np.int64)
    J, I = np.broadcast_arrays(J, I)
    x_out = (I < 0) | (I >= N)
    y_out = (J < 0) | (J >= N)
    Jc = np.where(x_out & y_out, np.clip(J, 0, N - 1), J)
    y_out = y_out & ~x_out
    fb = np.full(J.shape, face, dtype=np.int64)
    Jb = Jc.copy()
    Ib = I.copy()
    code = np.zeros(J.shape, dtype=np.int64)
    masks = {
        "E": x_out & (I >= N), "W": x_out & (I < 0),
        "N": y_out & (Jc >= N), "S": y_out & (Jc < 0),
    }
    for e, m in masks.items():
        if not m.any():
            continue
        f2, e2, rev, S = neighbour(face, e)
        if e == "E":
            d, k = I[m] - N, Jc[m]
        elif e == "W":
            d, k = -1 - I[m], Jc[m]
        elif e == "N":
            d, k = Jc[m] - N, I[m]
        else:
            d, k = -1 - Jc[m], I[m]
        if rev:
            k = N - 1 - k
        if e2 == "N":
            jb, ib = N - 1 - d, k
        elif e2 == "S":
            jb, ib = d, k
        elif e2 == "E":
            jb, ib = k, N - 1 - d
        else:
            jb, ib = k, d
        fb[m], Jb[m], Ib[m], code[m] = f2, jb, ib, perm_code(S)
    return fb, Jb, Ib, code


# ----------------------------------------------------------------------------
# tiles
# ----------------------------------------------------------------------------

def num_tiles(tiles_per_edge: int) -> int:
    """6 * p^2 (reference JAX-DevLab-Examples.py:40)."""
    return 6 * tiles_per_edge * tiles_per_edge


def tile_id(face, ty, tx, p):
    return face * p * p + ty * p + tx


def tile_coords(tile: int, p: int):
    face, r = divmod(tile, p * p)
    ty, tx = divmod(r, p)
    return face, ty, tx


def build_halo_table(N: int, p: int, g: int):
    """Flat gather table for filling every ghost cell of every tile.

    Returns a dict of int64 arrays, one entry per ghost cell:
      ``dst_tile, dst_j, dst_i`` -- ghost cell (padded tile coordinates),
      ``src_tile, src_j, src_i`` -- the interior cell it copies (padded coords),
      ``code`` -- vector-component code (0 for plain copy).
    Ghost cells of a tile that lie inside the face resolve to the neighbouring
    tile of the same face (code 0); those across a seam go through
    :func:`resolve`.
    """
    if N % p:
        raise ValueError(f"N = {N} is not divisible by tiles_per_edge = {p}")
    n = N // p
    if g < 1 or g > n:
        raise ValueError(f"halo width g = {g} must satisfy 1 <= g <= n = {n}")
    m = n + 2 * g
    jj, ii = np.meshgrid(np.arange(m), np.arange(m), indexing="ij")
    ghost = ~((jj >= g) & (jj < g + n) & (ii >= g) & (ii < g + n))
    gj, gi = jj[ghost], ii[ghost]
    out = {k: [] for k in ("dst_tile", "dst_j", "dst_i", "src_tile", "src_j", "src_i", "code")}
    for t in range(num_tiles(p)):
        face, ty, tx = tile_coords(t, p)
        J = ty * n + gj - g
        I = tx * n + gi - g
        fb, Jb, Ib, code = resolve(face, J, I, N)
        out["dst_tile"].append(np.full(gj.shape, t, dtype=np.int64))
        out["dst_j"].append(gj)
        out["dst_i"].append(gi)
        out["src_tile"].append(fb * p * p + (Jb // n) * p + (Ib // n))
        out["src_j"].append(Jb % n + g)
        out["src_i"].append(Ib % n + g)
        out["code"].append(code)
    return {k: np.concatenate(v) for k, v in out.items()}


def edge_push_descriptors(p: int):
    """Per (tile, edge) description of where a tile's *interior boundary strip* goes.

    Used by the fused stage kernel, which stores freshly computed boundary
    cells straight into the neighbour's ghost cells.  Returns an int32 array
    ``[T, 4, 4]`` with (dst_tile, dst_edge, reverse, code) where ``code`` maps
    this tile's components into the *destination* tile's axes.
    Edge order follows :data:`EDGES`.
    """
    T = num_tiles(p)
    desc = np.zeros((T, 4, 4), dtype=np.int32)
    for t in range(T):
        face, ty, tx = tile_coords(t, p)
        for e in EDGES:
            inside = {"N": ty < p - 1, "S": ty > 0, "E": tx < p - 1, "W": tx > 0}[e]
            if inside:
                dty = ty + (e == "N") - (e == "S")
                dtx = tx + (e == "E") - (e == "W")
                opposite = {"N": "S", "S": "N", "E": "W", "W": "E"}[e]
                desc[t, EDGE_ID[e]] = (tile_id(face, dty, dtx, p), EDGE_ID[opposite], 0, 0)
                continue
            fb, eb, rev, _ = neighbour(face, e)
            # components of *our* vectors seen from the neighbour: its own S for (fb, eb)
            _, _, _, S_b = neighbour(fb, eb)
            k = tx if e in ("N", "S") else ty          # tile index along the edge
            if rev:
                k = p - 1 - k
            if eb == "N":
                dty, dtx = p - 1, k
            elif eb == "S":
                dty, dtx = 0, k
            elif eb == "E":
                dty, dtx = k, p - 1
            else:
                dty, dtx = k, 0
            desc[t, EDGE_ID[e]] = (tile_id(fb, dty, dtx, p), EDGE_ID[eb], int(rev), perm_code(S_b))
    return desc
