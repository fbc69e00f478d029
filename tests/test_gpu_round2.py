"""Round-2 GPU tests: fused conservation sums, tiny tiles in the split stage, kernel bookkeeping, device hygiene.

All through the C ABI (ctypes), all against the CPU oracle or against another product path that is itself held to the
oracle elsewhere in this suite.
"""
import ctypes as C
import os

import numpy as np
import pytest

torch = pytest.importorskip("torch")

pytestmark = pytest.mark.gpu

from cubesphere import api, cases  # noqa: E402
from oracle import cs_oracle_np as orc  # noqa: E402


def _setup(name, N, p, dtype, g=2, **ctx_kw):
    case = cases.make_case(name, N)
    par = orc.Params(N, g, axis=case.axis, advect_only=case.advect_only, hs=case.hs)
    ref = orc.halo_fill_state(orc.pad(case.h, g), orc.pad(case.u1, g), orc.pad(case.u2, g), g)
    ctx = api.Context(N, tiles_per_edge=p, halo=g, dtype=dtype, advect_only=case.advect_only, axis=case.axis,
                      tangents=orc.centre_tangents(N), **ctx_kw)
    dev = [torch.from_numpy(orc.faces_to_tiles(x, p, g).astype(dtype)).cuda() for x in ref]
    if case.hs is not None:
        ctx.set_topography(torch.from_numpy(orc.faces_to_tiles(par.hs, p, g).astype(dtype)).cuda())
    return case, par, ref, ctx, dev


# ---------------------------------------------------------------------------------------------------------------
# K3 fused into the last stage of a step
# ---------------------------------------------------------------------------------------------------------------

@pytest.mark.parametrize("name,N,p", [("tc5", 96, 1), ("tc2", 48, 2), ("galewsky", 64, 1)])
def test_fused_step_diagnostics_fp64(built_library, name, N, p):
    """cs_step(..., diag_out): the sums come out of the last stage kernel.  fp64: mass 1e-13 / energy 1e-12 against
    the oracle and against the stand-alone reduction kernel on the same state; identical bits run to run."""
    case, par, ref, ctx, dev = _setup(name, N, p, np.float64)
    dt = cases.stable_dt(case, 0.5)
    diag = torch.full((2,), float("nan"), dtype=torch.float64, device="cuda")
    start = [x.clone() for x in dev]
    for _ in range(3):
        ctx.step(*dev, dt, diag=diag)
        ref = orc.step(par, ref, dt)
    assert ctx.stage_kernel_info()[0] == "march"
    got = diag.cpu().numpy()
    sep = ctx.diagnostics(*dev).cpu().numpy()
    want = np.array(orc.diagnostics(par, *ref))
    for other in (sep, want):
        assert abs(got[0] - other[0]) <= 1e-13 * abs(other[0]), (got, other)
        assert abs(got[1] - other[1]) <= 1e-12 * abs(other[1]), (got, other)
    # the same three steps again from the same start: the same bits (fixed summation order, no atomics on the sums)
    again = torch.zeros(2, dtype=torch.float64, device="cuda")
    for _ in range(3):
        ctx.step(*start, dt, diag=again)
    assert torch.equal(again, diag) and all(torch.equal(a, b) for a, b in zip(start, dev))
    ctx.close()


@pytest.mark.parametrize("entry", [False, True])
def test_fused_step_diagnostics_fp32(built_library, entry):
    """fp32 state: per-cell products are formed in fp32 (with the kernel's own fp32 metric) and accumulated in fp64.
    Against the all-fp64 reduction kernel on the same fp32 state the sums carry a constant relative offset of a few
    1e-8 (MUFU.RSQ in sqrt(G) is biased by a fraction of an ulp, and sqrt(G) ~ rsqrt^3): stated tolerance 1e-7.  The
    offset is a property of the grid, not of the state, so CHANGES of mass are seen much more finely: 2e-9."""
    case, par, ref, ctx, dev = _setup("tc5", 192, 1, np.float32, sums_of_entry_state=entry)
    dt = cases.stable_dt(case, 0.5)
    diag = torch.zeros(2, dtype=torch.float64, device="cuda")
    fused, sep = [], []
    for _ in range(4):
        if entry:
            sep.append(ctx.diagnostics(*dev).cpu().numpy())      # the state the step starts from
        ctx.step(*dev, dt, diag=diag)
        if not entry:
            sep.append(ctx.diagnostics(*dev).cpu().numpy())      # the new state
        fused.append(diag.cpu().numpy().copy())
    fused, sep = np.array(fused), np.array(sep)
    rel = np.abs(fused - sep) / np.abs(sep)
    print(f"fp32 fused (entry={entry}) vs fp64 reduction: mass {rel[:, 0].max():.2e}, energy {rel[:, 1].max():.2e}")
    assert rel.max() <= 1e-7, (fused, sep)
    drift_f = (fused[-1, 0] - fused[0, 0]) / fused[0, 0]
    drift_s = (sep[-1, 0] - sep[0, 0]) / sep[0, 0]
    assert abs(drift_f - drift_s) <= 2e-9, (drift_f, drift_s)
    ctx.close()


@pytest.mark.parametrize("name,N,p", [("tc5", 96, 1), ("tc2", 48, 2)])
def test_entry_state_sums_fp64(built_library, name, N, p):
    """sums_of_entry_state = 1: cs_step's diag_out is mass / energy of the state the step starts from, summed inside
    the FIRST stage launch.  fp64: 1e-13 / 1e-12 against the oracle's diagnostics of that state; same bits on a rerun."""
    case, par, ref, ctx, dev = _setup(name, N, p, np.float64, sums_of_entry_state=True)
    dt = cases.stable_dt(case, 0.5)
    diag = torch.full((2,), float("nan"), dtype=torch.float64, device="cuda")
    start = [x.clone() for x in dev]
    for _ in range(3):
        want = np.array(orc.diagnostics(par, *ref))
        ctx.step(*dev, dt, diag=diag)
        got = diag.cpu().numpy()
        assert abs(got[0] - want[0]) <= 1e-13 * abs(want[0]), (got, want)
        assert abs(got[1] - want[1]) <= 1e-12 * abs(want[1]), (got, want)
        ref = orc.step(par, ref, dt)
    again = torch.zeros(2, dtype=torch.float64, device="cuda")
    for _ in range(3):
        ctx.step(*start, dt, diag=again)
    assert torch.equal(again, diag) and all(torch.equal(a, b) for a, b in zip(start, dev))
    ctx.close()


def test_fused_diag_can_be_switched_off_and_stage_entry_agrees(built_library):
    """CS_FUSED_DIAG=0 (read when the context is created) sends diag_out through the reduction kernel; cs_rk_stage
    with diag_out gives the fused numbers too."""
    os.environ["CS_FUSED_DIAG"] = "0"
    try:
        case, par, ref, ctx0, dev0 = _setup("tc5", 48, 1, np.float64)
    finally:
        del os.environ["CS_FUSED_DIAG"]
    _, _, _, ctx1, dev1 = _setup("tc5", 48, 1, np.float64)
    dt = cases.stable_dt(case, 0.5)
    d0 = torch.zeros(2, dtype=torch.float64, device="cuda")
    d1 = torch.zeros(2, dtype=torch.float64, device="cuda")
    ctx0.step(*dev0, dt, diag=d0)
    ctx1.step(*dev1, dt, diag=d1)
    # the summing build of the stage kernel is a different instantiation: the compiler contracts a few multiply-adds
    # differently, so the states agree to rounding (each build is bit-reproducible run to run), not bit for bit
    for a, b in zip(dev0, dev1):
        scale = float(b.abs().max())
        worst = float((a - b).abs().max()) / scale
        print(f"state with / without fused sums: max difference {worst:.1e} of the field's scale")
        assert worst <= 1e-14
    a, b = d0.cpu().numpy(), d1.cpu().numpy()
    assert abs(a[0] - b[0]) <= 1e-13 * abs(a[0]) and abs(a[1] - b[1]) <= 1e-12 * abs(a[1])
    # one stage through cs_rk_stage
    out = [ctx1.new_field() for _ in range(3)]
    ds = torch.zeros(2, dtype=torch.float64, device="cuda")
    ctx1.rk_stage(dev1, dev1, out, 0.75, 0.25, dt, diag=ds)
    sep = ctx1.diagnostics(*out).cpu().numpy()
    got = ds.cpu().numpy()
    assert abs(got[0] - sep[0]) <= 1e-13 * abs(sep[0]) and abs(got[1] - sep[1]) <= 1e-12 * abs(sep[1])
    ctx0.close(), ctx1.close()


@pytest.mark.parametrize("of_input", [False, True])
@pytest.mark.parametrize("dtype,N,p", [(np.float64, 48, 2), (np.float64, 96, 1), (np.float32, 96, 2)])
def test_split_stage_sums_add_up(built_library, dtype, N, p, of_input):
    """cs_rk_stage_part_sums: inner-part sums (marching kernel) + rim-part sums (rim kernel) = the sums of the whole
    stage output (or input), which the stand-alone reduction kernel confirms."""
    case, par, ref, ctx, dev = _setup("tc5", N, p, dtype)
    dt = cases.stable_dt(case, 0.5)
    out = [ctx.new_field() for _ in range(3)]
    s4 = torch.full((4,), float("nan"), dtype=torch.float64, device="cuda")
    ctx.rk_stage(dev, dev, out, 0.75, 0.25, dt, part=0, diag=s4[:2], sums_of_input=of_input)
    ctx.rk_stage(dev, dev, out, 0.75, 0.25, dt, part=1, diag=s4[2:], sums_of_input=of_input)
    got = (s4[:2] + s4[2:]).cpu().numpy()
    want = ctx.diagnostics(*(dev if of_input else out)).cpu().numpy()
    tol = (1e-13, 1e-12) if dtype is np.float64 else (1e-7, 1e-7)
    assert abs(got[0] - want[0]) <= tol[0] * abs(want[0]) and abs(got[1] - want[1]) <= tol[1] * abs(want[1]), (got, want)
    rim_share = float(s4[2] / (s4[0] + s4[2]))
    n = N // p
    assert abs(rim_share - (1 - (n - 4) ** 2 / n ** 2)) < 0.05          # the rim is the rim: about 8 / n of the cells
    ctx.close()


# ---------------------------------------------------------------------------------------------------------------
# split stage on tiles with fewer than 2g cells per edge (every cell is a rim cell)
# ---------------------------------------------------------------------------------------------------------------

@pytest.mark.parametrize("N,p", [(12, 4), (8, 4), (12, 3)])          # n = 3, 2, 4 with g = 2
def test_split_stage_on_tiny_tiles(built_library, N, p):
    case, par, ref, ctx, dev = _setup("tc2", N, p, np.float64)
    dt = cases.stable_dt(case, 0.3)
    whole = [torch.full_like(x, float("nan")) for x in dev]
    split = [torch.full_like(x, float("nan")) for x in dev]
    ctx.rk_stage(dev, dev, whole, 0.75, 0.25, dt, part=2)
    ctx.rk_stage(dev, dev, split, 0.75, 0.25, dt, part=0)
    ctx.rk_stage(dev, dev, split, 0.75, 0.25, dt, part=1)
    g, n = 2, N // p
    c = slice(g, g + n)
    for a, b in zip(whole, split):
        assert not torch.isnan(b[:, c, c]).any(), "a cell was written by neither part"
        assert torch.equal(a[:, c, c], b[:, c, c])
    # and the whole stage is the oracle's
    want = orc.rk_stage(par, ref, ref, 0.75, 0.25, dt)
    got = orc.tiles_to_faces(whole[0].cpu().numpy(), p, g)
    cc = slice(g, g + N)
    assert np.allclose(got[:, cc, cc], want[0][:, cc, cc], rtol=1e-13, atol=0)
    ctx.close()


# ---------------------------------------------------------------------------------------------------------------
# bookkeeping and hygiene
# ---------------------------------------------------------------------------------------------------------------

def test_auto_fallback_to_the_slow_kernel_is_visible(built_library):
    """An odd row length cannot be vectorised: 'auto' takes the one-thread-per-cell kernel, and says so."""
    case, par, ref, ctx, dev = _setup("tc2", 25, 1, np.float64)     # m = 29
    assert ctx.stage_kernel_info() == ("auto", 0)
    dt = cases.stable_dt(case, 0.5)
    ctx.step(*dev, dt)
    name, falls = ctx.stage_kernel_info()
    assert name == "naive" and falls == 3
    ref = orc.step(par, ref, dt)
    c = slice(2, 27)
    assert np.allclose(dev[0].cpu().numpy()[:, c, c], ref[0][:, c, c], rtol=1e-13, atol=0)
    ctx.close()
    case, par, ref, ctx, dev = _setup("tc2", 24, 1, np.float64)
    ctx.step(*dev, cases.stable_dt(case, 0.5))
    assert ctx.stage_kernel_info() == ("march", 0)
    ctx.close()


def test_entry_points_restore_the_callers_device(built_library):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two visible GPUs")
    case, par, ref, ctx, dev = _setup("tc2", 24, 1, np.float64)       # context and state on cuda:0
    torch.cuda.synchronize(0)
    torch.cuda.set_device(1)                                           # the caller's thread works on another device
    try:
        null = C.c_void_p(None)
        lib = built_library
        rc = lib.cs_halo_fill_state(ctx._h, *(C.c_void_p(x.data_ptr()) for x in dev), null)
        assert rc == 0 and torch.cuda.current_device() == 1
        rc = lib.cs_step(ctx._h, *(C.c_void_p(x.data_ptr()) for x in dev), cases.stable_dt(case, 0.5), null, null)
        assert rc == 0 and torch.cuda.current_device() == 1
        # a context created and destroyed for device 0 while device 1 is current
        other = api.Context(24, halo=2, dtype=np.float64, device=0)
        assert torch.cuda.current_device() == 1
        other.close()
        ctx.close()
        assert torch.cuda.current_device() == 1
    finally:
        torch.cuda.set_device(0)
    torch.cuda.synchronize(0)
    ref = orc.step(par, ref, cases.stable_dt(case, 0.5))
    assert np.allclose(dev[0].cpu().numpy()[:, 2:26, 2:26], ref[0][:, 2:26, 2:26], rtol=1e-13, atol=0)


# ---------------------------------------------------------------------------------------------------------------
# a9: the vector halo with the geometric seam-rotation tables
# ---------------------------------------------------------------------------------------------------------------

@pytest.mark.parametrize("N,p", [(24, 1), (16, 2)])
def test_halo_with_geometric_rotation_tables(built_library, N, p):
    """geometry.seam_rotation_tables installed with cs_halo_set_rotation: the rotated ghosts of a solid-body wind are
    (1) bit-identical to the same two products and one sum in NumPy, and (2) the analytic wind's components in the
    ghost's own extended chart at the neighbour's cell centre, to 1e-12."""
    from cubesphere import geometry as geo
    g = 2
    n, m = N // p, N // p + 2 * g
    ctx = api.Context(N, tiles_per_edge=p, halo=g, dtype=np.float64)
    dst, src, code = ctx.halo_table()
    tab = {"dst_tile": dst // (m * m), "dst_j": (dst % (m * m)) // m, "dst_i": dst % m,
           "src_tile": src // (m * m), "src_j": (src % (m * m)) // m, "src_i": src % m, "code": code}
    R = geo.seam_rotation_tables(N, g, p, "contravariant", table=tab)
    ctx.set_rotation(R)
    # solid-body wind about a tilted axis, contravariant components on every tile's interior
    patch = geo.Patch(N, p)
    P = patch.points()
    w = np.array([0.3, -0.5, 0.8]) * 1e-5
    V = geo.EARTH_RADIUS * np.cross(np.broadcast_to(w, P.shape), P)
    u1, u2 = geo.contravariant_from_cartesian(V, N, patch=patch)
    u = np.full((6 * p * p, m, m), np.nan)
    v = np.full((6 * p * p, m, m), np.nan)
    u[:, g:g + n, g:g + n], v[:, g:g + n, g:g + n] = u1, u2
    du, dv = torch.from_numpy(u).cuda(), torch.from_numpy(v).cuda()
    ctx.halo_fill_vector(du, dv)
    gu, gv = du.cpu().numpy().reshape(-1), dv.cpu().numpy().reshape(-1)
    a, b = u.reshape(-1)[src], v.reshape(-1)[src]
    want_u = R[:, 0, 0] * a + R[:, 0, 1] * b              # NumPy: two products, one sum, no FMA
    want_v = R[:, 1, 0] * a + R[:, 1, 1] * b
    assert np.array_equal(gu[dst], want_u) and np.array_equal(gv[dst], want_v)
    assert not np.isnan(gu).any() and not np.isnan(gv).any()        # every ghost cell was written
    # analytic: the wind at the source cell's centre, expressed in the destination face's extended chart
    t = geo.centre_tangents(N)
    pp = p * p
    sf, df = tab["src_tile"] // pp, tab["dst_tile"] // pp
    r = tab["src_tile"] % pp
    X, Y = t[(r % p) * n + tab["src_i"] - g], t[(r // p) * n + tab["src_j"] - g]
    fr = np.asarray(geo.FACE_FRAMES, dtype=np.float64)[sf]
    Ps = fr[:, 0] + X[:, None] * fr[:, 1] + Y[:, None] * fr[:, 2]
    Ps /= np.linalg.norm(Ps, axis=-1, keepdims=True)
    Vs = geo.EARTH_RADIUS * np.cross(np.broadcast_to(w, Ps.shape), Ps)
    ea, eb, _, _ = geo.chart_basis(df, Ps)
    g11, g12, g22 = np.sum(ea * ea, -1), np.sum(ea * eb, -1), np.sum(eb * eb, -1)
    b1, b2 = np.sum(ea * Vs, -1), np.sum(eb * Vs, -1)
    det = g11 * g22 - g12 * g12
    au, av = (g22 * b1 - g12 * b2) / det, (g11 * b2 - g12 * b1) / det
    scale = max(np.abs(au).max(), np.abs(av).max())
    assert np.abs(gu[dst] - au).max() <= 1e-12 * scale and np.abs(gv[dst] - av).max() <= 1e-12 * scale
    ctx.set_rotation(None)
    ctx.close()


# ---------------------------------------------------------------------------------------------------------------
# one launch per step (step_march_kernel) against the one-stage-per-launch path
# ---------------------------------------------------------------------------------------------------------------

@pytest.mark.parametrize("name,N,p,dtype,rk", [("tc5", 48, 1, np.float64, "ssprk3"), ("tc2", 48, 2, np.float64, "ssprk3"),
                                                ("galewsky", 96, 1, np.float32, "ssprk3"), ("tc6", 40, 1, np.float64, "ssprk2"),
                                                ("tc1", 36, 3, np.float32, "ssprk3"), ("tc5", 192, 1, np.float64, "ssprk3")])
def test_step_kernel_matches_stage_launches(built_library, name, N, p, dtype, rk):
    """CS_STEP_KERNEL=1: cs_step in ONE cooperative launch (stages separated by grid barriers, edge ghosts pushed by the warps that
    computed the boundary cells, final table-driven fill) against the six-launch path: same interior to rounding, every
    ghost cell -- corner blocks included -- a bit-exact copy of its source, same entry-state sums, one kernel per step."""
    g = 2
    case = cases.make_case(name, N)
    kw = dict(tiles_per_edge=p, halo=g, dtype=dtype, rk=rk, advect_only=case.advect_only, axis=case.axis,
              tangents=orc.centre_tangents(N), sums_of_entry_state=True)
    ctx0 = api.Context(N, **kw)
    os.environ["CS_STEP_KERNEL"] = "1"                     # opt-in (read when the context is created)
    try:
        ctx1 = api.Context(N, **kw)
    finally:
        del os.environ["CS_STEP_KERNEL"]
    par = orc.Params(N, g, axis=case.axis, rk=rk, advect_only=case.advect_only, hs=case.hs)
    ref = orc.halo_fill_state(orc.pad(case.h, g), orc.pad(case.u1, g), orc.pad(case.u2, g), g)
    dev0 = [torch.from_numpy(orc.faces_to_tiles(x, p, g).astype(dtype)).cuda() for x in ref]
    dev1 = [x.clone() for x in dev0]
    if case.hs is not None:
        for c in (ctx0, ctx1):
            c.set_topography(torch.from_numpy(orc.faces_to_tiles(par.hs, p, g).astype(dtype)).cuda())
    dt = cases.stable_dt(case, 0.5)
    d0 = torch.zeros(2, dtype=torch.float64, device="cuda")
    d1 = torch.zeros(2, dtype=torch.float64, device="cuda")
    for _ in range(3):
        ctx0.step(*dev0, dt, diag=d0)
    l0 = ctx1.launches
    for _ in range(3):
        ctx1.step(*dev1, dt, diag=d1)
    assert ctx1.launches - l0 == 3, "expected one launch per step"
    tol = 1e-13 if dtype is np.float64 else 2e-6
    for a, b in zip(dev0, dev1):
        scale = float(a.abs().max())
        assert float((a - b).abs().max()) <= tol * scale
    # ghosts of the result: what the table-driven fill gives for this interior, bit for bit
    again = [x.clone() for x in dev1]
    ctx1.halo_fill_state(*again)
    assert all(torch.equal(a, b) for a, b in zip(again, dev1))
    a, b = d0.cpu().numpy(), d1.cpu().numpy()
    st = 1e-12 if dtype is np.float64 else 1e-7
    assert np.all(np.abs(a - b) <= st * np.abs(a)), (a, b)
    if dtype is np.float64:                                # and against the oracle
        for _ in range(3):
            ref = orc.step(par, ref, dt)
        got = orc.tiles_to_faces(dev1[0].cpu().numpy(), p, g)
        c = slice(g, g + N)
        assert np.linalg.norm(got[:, c, c] - ref[0][:, c, c]) <= 1e-13 * np.linalg.norm(ref[0][:, c, c])
    ctx0.close(), ctx1.close()


# ---------------------------------------------------------------------------------------------------------------
# second-order ghosts (cs_halo_set_interp): values at our own extended grid points
# ---------------------------------------------------------------------------------------------------------------

@pytest.mark.parametrize("N,p,dtype", [(24, 1, np.float64), (16, 2, np.float64), (24, 1, np.float32)])
def test_interpolated_halo_is_bit_exact(built_library, N, p, dtype):
    """The interpolating ghost fill against the same sums in NumPy, in the order the header states:
    ((w0 s0 + w1 s1) + w2 s2) + w3 s3, then two products and one sum per rotated component.  No FMA: same bits."""
    from cubesphere import geometry as geo
    g = 2
    n, m = N // p, N // p + 2 * g
    tabs = geo.seam_interp_tables(N, g, p)
    ctx = api.Context(N, tiles_per_edge=p, halo=g, dtype=dtype)
    ctx.set_halo_interpolation(tabs)
    rng = np.random.default_rng(3)
    f = [rng.standard_normal((6 * p * p, m, m)).astype(dtype) for _ in range(3)]
    dev = [torch.from_numpy(x.copy()).cuda() for x in f]
    ctx.halo_fill_state(*dev)
    w = tabs["w"].astype(dtype)
    R = tabs["rot"].astype(dtype)
    src, dst = tabs["src"], tabs["dst"]

    def interp(a):
        s = a.reshape(-1)[src]
        return ((w[:, 0] * s[:, 0] + w[:, 1] * s[:, 1]) + w[:, 2] * s[:, 2]) + w[:, 3] * s[:, 3]

    gh, ga, gb = (interp(x) for x in f)
    want = [x.copy() for x in f]
    want[0].reshape(-1)[dst] = gh
    want[1].reshape(-1)[dst] = R[:, 0, 0] * ga + R[:, 0, 1] * gb
    want[2].reshape(-1)[dst] = R[:, 1, 0] * ga + R[:, 1, 1] * gb
    for a, b in zip(dev, want):
        assert np.array_equal(a.cpu().numpy(), b)
    # the scalar entry point gives the same h; clearing the tables brings the plain copy back
    hs = ctx.halo_fill_scalar(torch.from_numpy(f[0].copy()).cuda())
    assert torch.equal(hs, dev[0])
    ctx.set_halo_interpolation(None)
    plain = [torch.from_numpy(x.copy()).cuda() for x in f]
    ctx.halo_fill_state(*plain)
    ref = orc.halo_fill_state(*(orc.tiles_to_faces(x.astype(np.float64), p, g) for x in f), g)
    assert np.array_equal(plain[0].cpu().numpy(), orc.faces_to_tiles(ref[0], p, g).astype(dtype))
    ctx.close()


@pytest.mark.parametrize("name", ["tc2", "tc5"])
def test_interpolated_halo_step_matches_the_oracle_study(built_library, name):
    """Full steps with interpolated ghosts and the extended metric against the oracle's statement of the same
    construction (oracle/seam_study.py, derived independently of geometry.seam_interp_tables): rel-L2(h) <= 1e-12."""
    from cubesphere import geometry as geo
    from oracle import seam_study
    N, g, steps = 32, 2, 20
    case = cases.make_case(name, N)
    halo = seam_study.SeamHalo(N, g, "interp")
    par = orc.Params(N, g, axis=case.axis, hs=None, tangents=halo.tangents, halo_fill=halo)
    if case.hs is not None:
        hs = orc.pad(case.hs, g)
        halo(hs, np.zeros_like(hs), np.zeros_like(hs))
        par.hs = hs
    ref = halo(orc.pad(case.h, g), orc.pad(case.u1, g), orc.pad(case.u2, g))
    ctx = api.Context(N, halo=g, dtype=np.float64, axis=case.axis, tangents=orc.centre_tangents(N))
    ctx.set_halo_interpolation(geo.seam_interp_tables(N, g))
    dev = [torch.from_numpy(orc.pad(x, g)).cuda() for x in (case.h, case.u1, case.u2)]
    if case.hs is not None:
        ctx.set_topography(torch.from_numpy(orc.pad(case.hs, g)).cuda())
    ctx.halo_fill_state(*dev)
    dt = cases.stable_dt(case, 0.5)
    for _ in range(steps):
        ctx.step(*dev, dt)
        ref = orc.step(par, ref, dt)
    c = slice(g, g + N)
    for a, b in zip(dev, ref):
        a = a.cpu().numpy()
        assert np.linalg.norm(a[:, c, c] - b[:, c, c]) <= 1e-12 * np.linalg.norm(b[:, c, c])
    ctx.close()


def test_interpolated_halo_is_second_order_on_the_gpu(built_library):
    """Williamson case 2 for half a day, limiter off: the L2 error of h falls by ~4 per doubling of the mesh with the
    interpolated ghosts (order > 1.8) and by ~2.3 with the plain copy (order < 1.5)."""
    from cubesphere import geometry as geo
    g = 2
    errs = {}
    for mode in ("copy", "interp"):
        for N in (24, 48):
            case = cases.make_case("tc2", N)
            ctx = api.Context(N, halo=g, dtype=np.float64, limiter="none", axis=case.axis)
            if mode == "interp":
                ctx.set_halo_interpolation(geo.seam_interp_tables(N, g))
            dev = [torch.from_numpy(orc.pad(x, g)).cuda() for x in (case.h, case.u1, case.u2)]
            ctx.halo_fill_state(*dev)
            dt = cases.stable_dt(case, 0.5)
            for _ in range(int(round(0.5 * 86400 / dt))):
                ctx.step(*dev, dt)
            c = slice(g, g + N)
            sg = geo.metric_terms(*np.meshgrid(geo.centre_tangents(N), geo.centre_tangents(N), indexing="xy"))["sg"]
            h = dev[0].cpu().numpy()[:, c, c]
            errs[mode, N] = np.sqrt(np.sum(sg * (h - case.h) ** 2) / np.sum(sg * case.h ** 2))
            ctx.close()
    order = {m: np.log2(errs[m, 24] / errs[m, 48]) for m in ("copy", "interp")}
    print("orders", order, errs)
    assert order["interp"] > 1.8 and order["copy"] < 1.5 and errs["interp", 48] < 0.2 * errs["copy", 48]


# ---------------------------------------------------------------------------------------------------------------
# f3: the reference's heat-diffusion demo on the scalar ghost fill
# ---------------------------------------------------------------------------------------------------------------

@pytest.mark.parametrize("N,p,dtype,tol", [(24, 1, np.float64, 1e-13), (32, 2, np.float64, 1e-13), (32, 1, np.float32, 5e-6)])
def test_diffusion_demo_matches_the_oracle(built_library, N, p, dtype, tol):
    """Checkerboard heat source on the top panel (deck slide 12), 100 explicit diffusion steps through
    cs_scalar_diffuse, against the oracle; four equal shares on the four neighbouring panels; heat conserved once the
    source is switched off."""
    g = 2
    par = orc.Params(N, g)
    src = orc.checkerboard_source(N, g)
    c = slice(g, g + N)
    kdt = 0.2 * par.delta ** 2 / max(par.gi11[c, c].max(), par.gi22[c, c].max())
    ctx = api.Context(N, tiles_per_edge=p, halo=g, dtype=dtype, tangents=orc.centre_tangents(N))
    a, b = ctx.new_field(), ctx.new_field()
    dsrc = torch.from_numpy(orc.faces_to_tiles(src, p, g).astype(dtype)).cuda()
    ref = np.zeros_like(src)
    for _ in range(100):
        ctx.scalar_diffuse(a, b, kdt, source=dsrc, dt_source=1.0)
        a, b = b, a
        ref = orc.diffuse(par, ref, kdt, src, 1.0)
    want = orc.faces_to_tiles(ref, p, g)
    assert np.linalg.norm(a.cpu().numpy() - want) <= tol * np.linalg.norm(want)      # every tile, ghosts included
    got = orc.tiles_to_faces(a.cpu().numpy().astype(np.float64), p, g)
    heat = lambda A: np.sum(par.sg[c, c] * A[:, c, c], axis=(1, 2))
    shares = heat(got)[1:5]
    assert np.ptp(shares) <= (1e-12 if dtype is np.float64 else 1e-5) * shares.mean()
    total0 = heat(got).sum()
    for _ in range(100):
        ctx.scalar_diffuse(a, b, kdt)
        a, b = b, a
    got = orc.tiles_to_faces(a.cpu().numpy().astype(np.float64), p, g)
    assert abs(heat(got).sum() - total0) <= (1e-13 if dtype is np.float64 else 1e-6) * total0
    with pytest.raises(ValueError):
        ctx.scalar_diffuse(a, a, kdt)
    ctx.close()


# ---------------------------------------------------------------------------------------------------------------
# f2: any order of the tiles along axis 0 (cs_config.tile_order)
# ---------------------------------------------------------------------------------------------------------------

@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_tile_order_is_only_a_relabelling(built_library, dtype):
    """A whole-sphere context whose axis 0 holds the tiles in a shuffled order: ghost fill, three steps, their sums and
    the diffusion demo give, tile for tile, the bits of the context with the plain order."""
    N, p, g = 48, 2, 2
    order = np.random.default_rng(5).permutation(6 * p * p).astype(np.int32)
    case, par, ref, ctx0, dev0 = _setup("tc5", N, p, dtype)
    ctx1 = api.Context(N, tiles_per_edge=p, halo=g, dtype=dtype, axis=case.axis, tangents=orc.centre_tangents(N), tile_order=order)
    perm = torch.from_numpy(order.astype(np.int64)).cuda()
    dev1 = [x[perm].contiguous() for x in dev0]
    ctx1.set_topography(torch.from_numpy(orc.faces_to_tiles(par.hs, p, g).astype(dtype)[order]).cuda())
    raw = [x.clone() for x in dev1]
    n = N // p
    for x in raw:                                           # poison the ghosts, fill, compare with the plain order's fill
        y = torch.full_like(x, float("nan"))
        y[:, g:g + n, g:g + n] = x[:, g:g + n, g:g + n]
        x.copy_(y)
    ctx1.halo_fill_state(*raw)
    assert all(torch.equal(a, b[perm]) for a, b in zip(raw, dev0))
    dt = cases.stable_dt(case, 0.5)
    d0 = torch.zeros(2, dtype=torch.float64, device="cuda")
    d1 = torch.zeros(2, dtype=torch.float64, device="cuda")
    for _ in range(3):
        ctx0.step(*dev0, dt, diag=d0)
        ctx1.step(*dev1, dt, diag=d1)
    assert all(torch.equal(a, b[perm]) for a, b in zip(dev1, dev0))
    a, b = d0.cpu().numpy(), d1.cpu().numpy()                # same cells, another summation order
    assert np.all(np.abs(a - b) <= 1e-13 * np.abs(a))
    assert np.all(np.abs(ctx1.diagnostics(*dev1).cpu().numpy() - ctx0.diagnostics(*dev0).cpu().numpy()) <= 1e-13 * np.abs(a))
    t0, t1 = ctx0.new_field(), ctx1.new_field()
    ctx0.scalar_diffuse(dev0[0], t0, 1e3)
    ctx1.scalar_diffuse(dev1[0], t1, 1e3)
    assert torch.equal(t1, t0[perm])
    with pytest.raises(ValueError):
        api.Context(N, tiles_per_edge=p, halo=g, tile_order=np.zeros(24, dtype=np.int32))
    ctx0.close(), ctx1.close()


def test_cli_diffusion_demo_and_second_order_flag(built_library, capsys):
    """python -m cubesphere.run --demo diffusion / --second-order: the two user-facing switches of this round."""
    from cubesphere import run
    run.main(["--demo", "diffusion", "--N", "24", "--steps", "40"])
    out = capsys.readouterr().out.strip().splitlines()
    shares = [float(x) for x in out[-1].split("(share of total):")[1].split("total")[0].split()]
    assert len(shares) == 6 and abs(sum(shares) - 1.0) < 1e-3
    assert max(shares[1:5]) - min(shares[1:5]) < 1e-4 and shares[0] > shares[1] > shares[5]
    totals = [float(l.split("total")[-1]) for l in out if "total" in l]
    assert abs(totals[-1] - totals[-3]) <= 1e-12 * totals[-1]          # no source in the second half: heat conserved
    run.main(["--case", "tc2", "--N", "24", "--hours", "2", "--second-order", "--limiter", "none"])
    out = capsys.readouterr().out
    lines = out.strip().splitlines()
    drift = float([l for l in lines if "mass drift =" in l][-1].split("mass drift =")[1].split()[0])
    assert 0 < abs(drift) < 1e-5                                        # truncation-level, not round-off, and small
    err = float([l for l in lines if "steady case" in l][-1].split("=")[1])
    assert 0 < err < 1e-3                                               # two hours of a steady flow at C24


@pytest.mark.parametrize("rk", ["euler", "ssprk2", "ssprk3"])
@pytest.mark.parametrize("entry", [False, True])
def test_step_sums_for_every_scheme(built_library, rk, entry):
    """The sums come out of the first stage (entry state) or the last one (new state) whatever the number of stages --
    also for the one-stage scheme, whose only stage is both."""
    N, g = 32, 2
    case = cases.make_case("tc5", N)
    par = orc.Params(N, g, axis=case.axis, rk=rk, hs=case.hs)
    ref = orc.halo_fill_state(orc.pad(case.h, g), orc.pad(case.u1, g), orc.pad(case.u2, g), g)
    ctx = api.Context(N, halo=g, dtype=np.float64, rk=rk, axis=case.axis, tangents=orc.centre_tangents(N), sums_of_entry_state=entry)
    ctx.set_topography(torch.from_numpy(par.hs).cuda())
    dev = [torch.from_numpy(x).cuda() for x in ref]
    dt = cases.stable_dt(case, 0.2)
    diag = torch.zeros(2, dtype=torch.float64, device="cuda")
    for _ in range(3):
        before = np.array(orc.diagnostics(par, *ref))
        ctx.step(*dev, dt, diag=diag)
        ref = orc.step(par, ref, dt)
        want = before if entry else np.array(orc.diagnostics(par, *ref))
        got = diag.cpu().numpy()
        assert abs(got[0] - want[0]) <= 1e-13 * abs(want[0]) and abs(got[1] - want[1]) <= 1e-12 * abs(want[1]), (rk, entry, got, want)
    c = slice(g, g + N)
    assert np.linalg.norm(dev[0].cpu().numpy()[:, c, c] - ref[0][:, c, c]) <= 1e-13 * np.linalg.norm(ref[0][:, c, c])
    ctx.close()


# ---------------------------------------------------------------------------------------------------------------
# the one-sided exchange with every rank's context in this process (pointers instead of IPC handles)
# ---------------------------------------------------------------------------------------------------------------

def _connect_in_process(ctxs, recv_counts):
    """recv_counts[r][q] = cells rank r receives from rank q."""
    for c in ctxs:
        c.halo_peer_alloc()
    ptrs = [c.halo_peer_pointers() for c in ctxs]
    world = len(ctxs)
    for r, c in enumerate(ctxs):
        c.halo_peer_connect_pointers(r, ptrs, [sum(recv_counts[q][:r]) for q in range(world)])


@pytest.mark.parametrize("second_order", [False, True])
@pytest.mark.parametrize("p,world,tile_map", [(1, 2, "block"), (2, 4, "locality"), (2, 8, "locality"), (2, 8, "block")])
def test_one_sided_exchange_emulated_on_one_gpu(built_library, p, world, tile_map, second_order):
    """halo_push / halo_pull (the kernels of the multi-GPU exchange) with all ranks on this GPU: pushes store into the other
    contexts' receive buffers and raise their flags, pulls wait for the flags and fill the ghosts.  Copy ghosts: bit-exact
    against the oracle's fill.  Second-order ghosts: raw cells travel, the receiver interpolates -- bit-identical to the
    whole-sphere context with the same tables.  Two fills in a row (both buffer parities), then three steps."""
    from cubesphere import geometry as geo, partition
    from cubesphere.sharded import RK_PAIRS
    N, g = 48, 2
    n, m = N // p, N // p + 2 * g
    case = cases.make_case("tc2", N, alpha0=0.6)
    order = np.arange(6 * p * p) if tile_map == "block" else partition.locality_order(p, world)
    ranges = partition.block_ranges(6 * p * p, world)
    tang = orc.centre_tangents(N)
    whole = api.Context(N, tiles_per_edge=p, halo=g, axis=case.axis, tangents=tang)
    if second_order:
        whole.set_halo_interpolation(geo.seam_interp_tables(N, g, p))
    ctxs, recv_counts = [], []
    for r, (b, c) in enumerate(ranges):
        ctx = api.Context(N, tiles_per_edge=p, halo=g, axis=case.axis, tangents=tang, tile_begin=b, tile_count=c, tile_order=order)
        if second_order:
            plan = partition.InterpExchangePlan(N, p, g, r, ranges, order)
            ctx.halo_interp_sharded(plan)
            recv_counts.append(plan.recv_counts)
        else:
            _, recv = ctx.halo_plan([x for x, _ in ranges], [x for _, x in ranges])
            recv_counts.append(recv)
        ctxs.append(ctx)
    _connect_in_process(ctxs, recv_counts)
    rng = np.random.default_rng(2)
    for trial in range(2):
        tiles = [np.ascontiguousarray(rng.standard_normal((6 * p * p, m, m))) for _ in range(3)]
        ref = [torch.from_numpy(x.copy()).cuda() for x in tiles]
        whole.halo_fill_state(*ref)
        shards = []
        for r, (b, c) in enumerate(ranges):
            sh = []
            for x in tiles:
                y = np.full((c, m, m), np.nan)
                y[:, g:g + n, g:g + n] = x[order[b:b + c]][:, g:g + n, g:g + n]
                sh.append(torch.from_numpy(y).cuda())
            shards.append(sh)
        for ctx, sh in zip(ctxs, shards):
            ctx.halo_fill_state(*sh)
        for ctx, sh in zip(ctxs, shards):
            ctx.halo_push(*sh)
        for ctx, sh in zip(ctxs, shards):
            ctx.halo_pull(*sh)
        idx = torch.from_numpy(order.astype(np.int64)).cuda()
        for (b, c), sh in zip(ranges, shards):
            for got, want in zip(sh, ref):
                assert torch.equal(got, want[idx[b:b + c]])
    assert all(c.halo_peer_timeouts() == 0 for c in ctxs)
    # three steps of the steady case, stage by stage as the sharded stepper runs them
    par_ref = [torch.from_numpy(orc.faces_to_tiles(orc.pad(x, g), p, g)).cuda() for x in (case.h, case.u1, case.u2)]
    whole.halo_fill_state(*par_ref)
    state = [[x[idx[b:b + c]].contiguous() for x in par_ref] for b, c in ranges]
    qa = [[c.new_field() for _ in range(3)] for c in ctxs]
    qb = [[c.new_field() for _ in range(3)] for c in ctxs]
    dt = cases.stable_dt(case, 0.5)
    for _ in range(3):
        whole.step(*par_ref, dt)
        cur = state
        for k, (a, bb) in enumerate(RK_PAIRS["ssprk3"]):
            out = state if k == 2 else (qa if k == 0 else qb)
            for r, ctx in enumerate(ctxs):
                ctx.rk_stage(state[r], cur[r], out[r], a, bb, dt, part=1)
                ctx.rk_stage(state[r], cur[r], out[r], a, bb, dt, part=0)
                ctx.halo_fill_state(*out[r])
            for r, ctx in enumerate(ctxs):
                ctx.halo_push(*out[r])
            for r, ctx in enumerate(ctxs):
                ctx.halo_pull(*out[r])
            cur = out
    for (b, c), sh in zip(ranges, state):
        for got, want in zip(sh, par_ref):
            w = want[idx[b:b + c]]
            assert float((got - w)[:, g:g + n, g:g + n].abs().max()) <= 1e-13 * float(w.abs().max())
    torch.cuda.synchronize()
    for c in ctxs:
        c.close()
    whole.close()


def test_one_sided_pull_gives_up_and_says_so(built_library):
    """A pull whose peer never pushes must not hang the GPU: it gives up after `peer_timeout_ms`, counts the event,
    and from then on no pull waits at all (the job drains instead of paying the time-out on every fill)."""
    import time
    N, g, p, world = 24, 2, 1, 2
    from cubesphere import partition
    ranges = partition.block_ranges(6, world)
    ctxs, recv_counts = [], []
    for b, c in ranges:
        ctx = api.Context(N, tiles_per_edge=p, halo=g, tile_begin=b, tile_count=c)
        _, recv = ctx.halo_plan([x for x, _ in ranges], [x for _, x in ranges])
        recv_counts.append(recv)
        ctx.set_option("peer_timeout_ms", 200)
        ctxs.append(ctx)
    _connect_in_process(ctxs, recv_counts)
    m = N + 2 * g
    sh = [[torch.zeros((c, m, m), dtype=torch.float64, device="cuda") for _ in range(3)] for _, c in ranges]
    assert ctxs[0].halo_peer_timeouts() == 0
    ctxs[0].halo_push(*sh[0])                    # rank 0 does its part, rank 1 never pushes
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    ctxs[0].halo_pull(*sh[0])
    torch.cuda.synchronize()
    waited = time.perf_counter() - t0
    assert 0.1 < waited < 2.0, waited            # about the 200 ms asked for, not the default and not for ever
    assert ctxs[0].halo_peer_timeouts() >= 1
    t0 = time.perf_counter()
    for _ in range(5):                           # sticky: later pulls do not wait
        ctxs[0].halo_push(*sh[0])
        ctxs[0].halo_pull(*sh[0])
    torch.cuda.synchronize()
    assert time.perf_counter() - t0 < 0.1
    # the healthy direction still works: rank 1 has rank 0's pushes
    ctxs[1].halo_push(*sh[1])
    ctxs[1].halo_pull(*sh[1])
    torch.cuda.synchronize()
    assert ctxs[1].halo_peer_timeouts() == 0
    for c in ctxs:
        c.close()
