"""Host logic: cube topology against the reference's own table, tiles, sharding validation."""
import numpy as np
import pytest

from cubesphere import api, partition, topology
from oracle import cs_oracle_np as orc

# Golden data from the reference: the 12 seam swaps in 4 stages with their op flags
# (/root/reference/JAX-DevLab-Examples.py:114-139; identical on slide 9 of the deck).
REFERENCE_SCHEDULE = (
    (((0, "N"), (1, "N"), "R"), ((3, "E"), (4, "W"), "N"), ((2, "S"), (5, "E"), "TR")),
    (((0, "E"), (4, "N"), "T"), ((2, "E"), (3, "W"), "N"), ((1, "S"), (5, "N"), "N")),
    (((0, "W"), (2, "N"), "TR"), ((1, "W"), (4, "E"), "N"), ((3, "S"), (5, "S"), "R")),
    (((0, "S"), (3, "N"), "N"), ((1, "E"), (2, "W"), "N"), ((4, "S"), (5, "W"), "T")),
)


def test_schedule_equals_reference_table():
    assert topology.create_communication_schedule() == REFERENCE_SCHEDULE
    api.validate_schedule(REFERENCE_SCHEDULE)


def test_oracle_adjacency_equals_reference_table():
    adj = orc.adjacency()
    for stage in REFERENCE_SCHEDULE:
        faces = []
        for (fa, ea), (fb, eb), ops in stage:
            assert adj[(fa, ea)][:2] == (fb, eb) and adj[(fb, eb)][:2] == (fa, ea)
            assert orc.op_flag(fa, ea) == ops == orc.op_flag(fb, eb)
            faces += [fa, fb]
        assert sorted(faces) == list(range(6))          # perfect matching: no face twice in a stage


def test_frames_right_handed_and_seam_maps_inverse():
    for n, x, y in topology.FACE_FRAMES:
        assert np.array_equal(np.cross(x, y), n)
    for f in range(6):
        for e in topology.EDGES:
            fb, eb, rev, S = topology.neighbour(f, e)
            f2, e2, rev2, S2 = topology.neighbour(fb, eb)
            assert (f2, e2, rev2) == (f, e, rev)
            assert np.array_equal(S @ S2, np.eye(2, dtype=np.int64))    # there and back is the identity
            assert abs(np.linalg.det(S)) == 1


def test_apply_operations_matches_reference_semantics():
    d = np.arange(5.0)
    assert np.array_equal(topology.apply_operations(d, "N"), d)
    assert np.array_equal(topology.apply_operations(d, "T"), d)
    assert np.array_equal(topology.apply_operations(d, "R"), d[::-1])
    assert np.array_equal(topology.apply_operations(d, "TR"), d[::-1])
    with pytest.raises(ValueError, match="seam operation must be one of"):
        topology.apply_operations(d, "X")


def test_validate_schedule_rejects_bad_tables():
    bad = list(map(list, REFERENCE_SCHEDULE))
    bad[0][0] = ((0, "N"), (1, "N"), "N")               # wrong flag
    with pytest.raises(ValueError):
        api.validate_schedule(bad)
    bad[0][0] = ((0, "N"), (2, "N"), "R")               # not a seam
    with pytest.raises(ValueError):
        api.validate_schedule(bad)
    with pytest.raises(ValueError):
        api.validate_schedule(REFERENCE_SCHEDULE[:3])   # 9 of 12 seams
    with pytest.raises(ValueError, match="seam operation must be one of"):
        api.validate_schedule(((((0, "N"), (1, "N"), "Q"),),))


@pytest.mark.parametrize("N,p,g", [(8, 1, 1), (8, 1, 2), (8, 2, 2), (12, 3, 2), (12, 2, 3)])
def test_gather_table_equals_oracle_slicing(N, p, g):
    """Two independent statements of the halo fill: per-cell gather table vs whole-strip slicing."""
    rng = np.random.default_rng(1)
    M = N + 2 * g
    h, u, v = (rng.standard_normal((6, M, M)) for _ in range(3))
    ho, uo, vo = orc.halo_fill_state(h, u, v, g)
    tab = topology.build_halo_table(N, p, g)
    n = N // p

    def tiles(a):
        t = orc.faces_to_tiles(a, p, g)
        w = np.full_like(t, np.nan)
        w[:, g:g + n, g:g + n] = t[:, g:g + n, g:g + n]
        return w

    ht, ut, vt = tiles(h), tiles(u), tiles(v)
    s = (tab["src_tile"], tab["src_j"], tab["src_i"])
    d = (tab["dst_tile"], tab["dst_j"], tab["dst_i"])
    ht[d] = ht[s]
    a, b = partition.apply_code(ut[s], vt[s], tab["code"])
    ut[d], vt[d] = a, b
    for got, ref in ((ht, ho), (ut, uo), (vt, vo)):
        assert np.array_equal(got, orc.faces_to_tiles(ref, p, g))


def test_reference_helpers_numpy():
    N, g = 6, 1
    a = np.arange((N + 2) ** 2, dtype=np.float64).reshape(N + 2, N + 2)
    assert np.array_equal(api.extract_boundary_data(a, "N", N), a[N, 1:N + 1])
    assert np.array_equal(api.extract_boundary_data(a, "S", N), a[1, 1:N + 1])
    assert np.array_equal(api.extract_boundary_data(a, "E", N), a[1:N + 1, N])
    assert np.array_equal(api.extract_boundary_data(a, "W", N), a[1:N + 1, 1])
    line = np.full(N, -1.0)
    for e, sl in (("N", (N + 1, slice(1, N + 1))), ("S", (0, slice(1, N + 1))),
                  ("E", (slice(1, N + 1), N + 1)), ("W", (slice(1, N + 1), 0))):
        out = api.set_ghost_data(a, e, line, N)
        assert np.array_equal(out[sl], line) and out is not a
        out[sl] = a[sl]
        assert np.array_equal(out, a)                    # nothing else changed
    with pytest.raises(ValueError):
        api.extract_boundary_data(a, "Q", N)


def test_setup_sharding_validation():
    mesh, sh = api.setup_sharding({"parallelization": {"tiles_per_edge": 1, "num_devices": 6, "device_type": "cpu"}})
    assert mesh.size == 6 and mesh.axis_names == ("tiles",) and sh.tiles_per_device == 1
    mesh, sh = api.setup_sharding({"parallelization": {}})                       # reference defaults: cpu, 6, 1
    assert mesh.size == 6
    mesh, sh = api.setup_sharding({"parallelization": {"tiles_per_edge": 2, "num_devices": 8}})
    assert sh.num_tiles == 24 and sh.tile_range(3) == (9, 3) and sh.owner(10) == 3
    with pytest.raises(ValueError, match="at least one tile"):
        api.setup_sharding({"parallelization": {"tiles_per_edge": 1, "num_devices": 7}})
    with pytest.raises(ValueError, match="cannot be split evenly"):
        api.setup_sharding({"parallelization": {"tiles_per_edge": 1, "num_devices": 4}})
    with pytest.raises(ValueError):
        api.setup_sharding({"parallelization": {"device_type": "tpu"}})


def test_push_descriptors_consistent_with_table():
    """The (tile, edge) -> (tile, edge, reverse, code) descriptors agree with the gather table."""
    N, p, g = 12, 2, 2
    n = N // p
    desc = topology.edge_push_descriptors(p)
    tab = topology.build_halo_table(N, p, g)
    key = {(int(t), int(j), int(i)): (int(st), int(sj), int(si), int(c)) for t, j, i, st, sj, si, c in
           zip(tab["dst_tile"], tab["dst_j"], tab["dst_i"], tab["src_tile"], tab["src_j"], tab["src_i"], tab["code"])}
    for t in range(topology.num_tiles(p)):
        for e, name in enumerate(topology.EDGES):
            dt_, de, rev, code = (int(x) for x in desc[t, e])
            for d in range(g):
                for k in range(n):
                    sj, si = {"N": (g + n - 1 - d, g + k), "S": (g + d, g + k),
                              "E": (g + k, g + n - 1 - d), "W": (g + k, g + d)}[name]
                    kk = n - 1 - k if rev else k
                    dj, di = {0: (g + n + d, g + kk), 1: (g - 1 - d, g + kk),
                              2: (g + kk, g + n + d), 3: (g + kk, g - 1 - d)}[de]
                    assert key[(dt_, dj, di)] == (t, sj, si, code)


def test_load_config_defaults_and_validation(tmp_path):
    from cubesphere import config
    cfg = config.load_config()
    assert cfg["parallelization"] == {"tiles_per_edge": 1, "num_devices": 6, "device_type": "cpu"}   # PY:22-24
    cfg = config.load_config({"parallelization": {"tiles_per_edge": 2, "num_devices": 8}, "extra": {"a": 1}})
    assert cfg["parallelization"]["tiles_per_edge"] == 2 and cfg["extra"] == {"a": 1}
    mesh, sh = api.setup_sharding(cfg)
    assert mesh.size == 8 and sh.tiles_per_device == 3
    with pytest.raises(ValueError, match="unknown key"):
        config.load_config({"parallelization": {"tile_per_edge": 2}})
    f = tmp_path / "config.yaml"
    f.write_text("parallelization:\n  tiles_per_edge: 1\n  num_devices: 3\n  device_type: 'cpu'\nsolver:\n  N: 96\n")
    cfg = config.load_config(f)
    assert cfg["parallelization"]["num_devices"] == 3 and cfg["solver"]["N"] == 96 and cfg["solver"]["halo"] == 2


@pytest.mark.parametrize("p", [1, 2, 3, 4])
def test_tile_schedule_is_a_proper_edge_colouring(p):
    sched = topology.create_tile_schedule(p)
    if p == 1:
        assert sched == REFERENCE_SCHEDULE
    assert len(sched) <= 5
    seen = set()
    for stage in sched:
        tiles = [t for (ta, _), (tb, _), _ in stage for t in (ta, tb)]
        assert len(tiles) == len(set(tiles))                       # no tile twice in a stage
        for (ta, ea), (tb, eb), ops in stage:
            key = frozenset(((ta, ea), (tb, eb)))
            assert key not in seen and ops in ("N", "T", "R", "TR")
            seen.add(key)
    assert len(seen) == 12 * p * p                                  # every tile edge pair exactly once


def test_state_files_round_trip(tmp_path):
    from cubesphere import io
    rng = np.random.default_rng(0)
    state = {k: rng.standard_normal((24, 10, 10)).astype(np.float32) for k in ("h", "u", "v")}
    f = io.save_state(tmp_path / "hist" / "s0001.npz", state, N=12, tiles_per_edge=2, halo=2, step=7, time=123.5, case="tc2")
    got, meta = io.load_state(f)
    assert all(np.array_equal(got[k], state[k]) for k in state)
    assert (meta["N"], meta["tiles_per_edge"], meta["halo"], meta["step"], meta["time"], meta["case"]) == (12, 2, 2, 7, 123.5, "tc2")
    with pytest.raises(ValueError):
        io.save_state(tmp_path / "bad.npz", state, N=16, halo=2)


# ---- property tests (hypothesis) -------------------------------------------------------------------
from hypothesis import given, settings, strategies as st  # noqa: E402


@settings(max_examples=200, deadline=None)
@given(face=st.integers(0, 5), N=st.integers(4, 40), g=st.integers(1, 3), data=st.data())
def test_resolve_lands_inside_the_neighbour_and_mirrors_back(face, N, g, data):
    """A ghost cell across one edge resolves to an interior cell of the neighbour within depth g of the shared edge;
    the neighbour's ghost at the mirrored position resolves back to our boundary cell with the inverse code."""
    e = data.draw(st.sampled_from(topology.EDGES))
    d = data.draw(st.integers(0, g - 1))
    k = data.draw(st.integers(0, N - 1))
    J, I = {"N": (N + d, k), "S": (-1 - d, k), "E": (k, N + d), "W": (k, -1 - d)}[e]
    fb, Jb, Ib, code = (int(x) for x in topology.resolve(face, J, I, N))
    f2, e2, rev, S = topology.neighbour(face, e)
    assert fb == f2 and 0 <= Jb < N and 0 <= Ib < N and code == topology.perm_code(S)
    depth = {"N": N - 1 - Jb, "S": Jb, "E": N - 1 - Ib, "W": Ib}[e2]
    assert depth == d                                              # ghost depth d <- interior depth d
    # our interior cell at depth d next to the edge, seen from the neighbour's ghost zone
    along = Ib if e2 in ("N", "S") else Jb
    Jg, Ig = {"N": (N + d, along), "S": (-1 - d, along), "E": (along, N + d), "W": (along, -1 - d)}[e2]
    f0, J0, I0, code_back = (int(x) for x in topology.resolve(f2, Jg, Ig, N))
    assert f0 == face
    ours = {"N": (N - 1 - d, k), "S": (d, k), "E": (k, N - 1 - d), "W": (k, d)}[e]
    assert (J0, I0) == ours
    u, v = np.array([3.0]), np.array([-5.0])
    a, b = partition.apply_code(*partition.apply_code(u, v, np.array([code])), np.array([code_back]))
    assert a[0] == u[0] and b[0] == v[0]                            # there and back: the identity on vectors


@settings(max_examples=60, deadline=None)
@given(p=st.integers(1, 4), world=st.integers(1, 24))
def test_block_ranges_partition_or_reject(p, world):
    T = topology.num_tiles(p)
    if world > T or T % world:
        with pytest.raises(ValueError):
            partition.block_ranges(T, world)
        return
    ranges = partition.block_ranges(T, world)
    assert len(ranges) == world and ranges[0][0] == 0 and sum(c for _, c in ranges) == T
    assert all(b2 == b1 + c1 for (b1, c1), (b2, _) in zip(ranges, ranges[1:]))
    assert len({c for _, c in ranges}) == 1


@settings(max_examples=25, deadline=None)
@given(p=st.integers(1, 3), n=st.integers(2, 5), g=st.integers(1, 2))
def test_halo_table_covers_every_ghost_cell_once(p, n, g):
    N = n * p
    if g > n:
        return
    tab = topology.build_halo_table(N, p, g)
    m = n + 2 * g
    T = topology.num_tiles(p)
    ghosts = T * (m * m - n * n)
    assert len(tab["dst_tile"]) == ghosts
    key = tab["dst_tile"] * m * m + tab["dst_j"] * m + tab["dst_i"]
    assert len(np.unique(key)) == ghosts                            # one writer per ghost cell
    sj, si = tab["src_j"], tab["src_i"]
    assert np.all((sj >= g) & (sj < g + n) & (si >= g) & (si < g + n))   # sources are interior cells only
    assert set(np.unique(tab["code"])) <= set(range(8))
