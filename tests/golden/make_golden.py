#!/usr/bin/env python
"""Generates the fixtures in this directory.  Run from the repository root:  python tests/golden/make_golden.py

What they pin, and what they cannot.  The reference ships no tests, no fixtures and no runnable solver
(and JAX is not installed), so there is nothing upstream to generate vectors from.  The one piece of
reference DATA, the seam table of /root/reference/JAX-DevLab-Examples.py:114-139, is transcribed in
`reference_schedule.json` (facts, not code).  Everything else here is output of the clean-room NumPy oracle
(oracle/cs_oracle_np.py) frozen at the commit that introduced this file: the tests compare the NumPy oracle,
its C restatement and the CUDA path against these frozen numbers, so a later change of any of the three
definitions shows up as a diff against committed data rather than going unnoticed.

  halo_c8_g2.npz    random (h, u, v) on 6 x 12 x 12 faces and their ghost-filled versions (bit-exact targets)
  halo_c12_p2.npz   the same in the tile layout, tiles_per_edge = 2
  step_tc2_c16.npz  Williamson case 2 (alpha = pi/4), C16: state after 5 SSP-RK3 / MC steps, mass, energy
  step_tc5_c16.npz  Williamson case 5 (mountain), C16: the same, with the topography
  step_tc1_c16.npz  Williamson case 1 (advection only, alpha = pi/4), C16, SSP-RK2 / minmod
  diffuse_c12.npz   (round 2) the heat-diffusion demo: checkerboard source on face 0, C12, field after 30 steps
  interp_c12.npz    (round 2) random (h, u, v) on C12 with the edge ghosts built by 1-D interpolation + rotation
                    (oracle/seam_study.py, mode "interp"); corner blocks are not part of that construction
Run with an argument (`python tests/golden/make_golden.py round2`) to write only the round-2 files.
"""
import json
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "sharding-the-sphere-fall-2025-jax-devlab-examples_b200"))
from cubesphere import cases  # noqa: E402
from oracle import cs_oracle_np as orc  # noqa: E402

HERE = Path(__file__).resolve().parent

REFERENCE_SCHEDULE = [      # stage -> [[face, edge], [face, edge], operation]
    [[[0, "N"], [1, "N"], "R"], [[3, "E"], [4, "W"], "N"], [[2, "S"], [5, "E"], "TR"]],
    [[[0, "E"], [4, "N"], "T"], [[2, "E"], [3, "W"], "N"], [[1, "S"], [5, "N"], "N"]],
    [[[0, "W"], [2, "N"], "TR"], [[1, "W"], [4, "E"], "N"], [[3, "S"], [5, "S"], "R"]],
    [[[0, "S"], [3, "N"], "N"], [[1, "E"], [2, "W"], "N"], [[4, "S"], [5, "W"], "T"]],
]


def halo_fixture(name, N, g, p):
    rng = np.random.default_rng(20251118)
    M = N + 2 * g
    h, u, v = (rng.standard_normal((6, M, M)) for _ in range(3))
    ho, uo, vo = orc.halo_fill_state(h, u, v, g)
    so = orc.halo_fill_scalar(h, g)
    assert np.array_equal(so, ho)
    arrs = {"h_in": h, "u_in": u, "v_in": v, "h_out": ho, "u_out": uo, "v_out": vo}
    if p > 1:
        arrs = {k: orc.faces_to_tiles(a, p, g) for k, a in arrs.items()}
    np.savez_compressed(HERE / name, N=N, g=g, p=p, **arrs)


def step_fixture(name, case_name, N, nsteps, rk, limiter, **kw):
    g = 2
    case = cases.make_case(case_name, N, **kw)
    par = orc.Params(N, g, axis=case.axis, advect_only=case.advect_only, hs=case.hs, rk=rk, limiter=limiter)
    state = orc.halo_fill_state(orc.pad(case.h, g), orc.pad(case.u1, g), orc.pad(case.u2, g), g)
    dt = cases.stable_dt(case)
    d0 = orc.diagnostics(par, *state)
    out = state
    for _ in range(nsteps):
        out = orc.step(par, out, dt)
    d1 = orc.diagnostics(par, *out)
    np.savez_compressed(HERE / name, N=N, g=g, nsteps=nsteps, dt=dt, axis=np.asarray(case.axis), rk=rk, limiter=limiter,
                        advect_only=bool(case.advect_only), hs=par.hs, h0=state[0], u0=state[1], v0=state[2],
                        h1=out[0], u1=out[1], v1=out[2], diag0=np.asarray(d0), diag1=np.asarray(d1))


def diffuse_fixture(name, N, nsteps):
    g = 2
    par = orc.Params(N, g)
    src = orc.checkerboard_source(N, g)
    c = slice(g, g + N)
    kdt = 0.2 * par.delta ** 2 / max(par.gi11[c, c].max(), par.gi22[c, c].max())
    T = np.zeros_like(src)
    for _ in range(nsteps):
        T = orc.diffuse(par, T, kdt, src, 1.0)
    np.savez_compressed(HERE / name, N=N, g=g, nsteps=nsteps, kdt=kdt, source=src, T=T)


def interp_fixture(name, N):
    from oracle import seam_study
    g = 2
    rng = np.random.default_rng(20251118)
    M = N + 2 * g
    h, u, v = (rng.standard_normal((6, M, M)) for _ in range(3))
    halo = seam_study.SeamHalo(N, g, "interp")
    ho, uo, vo = halo(h.copy(), u.copy(), v.copy())
    np.savez_compressed(HERE / name, N=N, g=g, h_in=h, u_in=u, v_in=v, h_out=ho, u_out=uo, v_out=vo, dst=halo.dst,
                        tangents=halo.tangents)


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "round2":
        diffuse_fixture("diffuse_c12.npz", 12, 30)
        interp_fixture("interp_c12.npz", 12)
        sys.exit(0)
    (HERE / "reference_schedule.json").write_text(json.dumps(REFERENCE_SCHEDULE) + "\n")
    halo_fixture("halo_c8_g2.npz", 8, 2, 1)
    halo_fixture("halo_c12_p2.npz", 12, 2, 2)
    step_fixture("step_tc2_c16.npz", "tc2", 16, 5, "ssprk3", "mc", alpha0=np.pi / 4)
    step_fixture("step_tc5_c16.npz", "tc5", 16, 5, "ssprk3", "mc")
    step_fixture("step_tc1_c16.npz", "tc1", 16, 5, "ssprk2", "minmod", alpha0=np.pi / 4)
    diffuse_fixture("diffuse_c12.npz", 12, 30)
    interp_fixture("interp_c12.npz", 12)
    for f in sorted(HERE.glob("*.npz")):
        print(f.name, f.stat().st_size, "bytes")
