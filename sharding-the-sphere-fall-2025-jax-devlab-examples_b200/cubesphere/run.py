"""Command-line driver: integrate one of the test cases on one GPU and print the conservation diagnostics.

    python -m cubesphere.run --case tc5 --N 384 --hours 24 --dtype f64 [--restart out.npz]
    python -m cubesphere.run --resume out.npz --hours 24 [--restart out2.npz]      # carry on from a restart file
    python -m cubesphere.run --case tc6 --N 192 --hours 48 --history run.zarr --history-every 100   # zarr v2 history
    python -m cubesphere.run --resume run.zarr --hours 24                          # ... and carry on from its last record
    python -m cubesphere.run --case tc2 --N 96 --second-order                      # interpolated ghosts
    python -m cubesphere.run --demo diffusion --N 96 --steps 2000                  # the reference's heat-diffusion demo

A thin example of the public API (cases -> Context -> step -> diagnostics -> restart file); the hot path is
entirely inside the C ABI calls.
"""
from __future__ import annotations

import argparse
import time

import numpy as np


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    ap.add_argument("--case", default="tc2")
    ap.add_argument("--N", type=int, default=96)
    ap.add_argument("--dtype", default="f64", choices=["f32", "f64"])
    ap.add_argument("--hours", type=float, default=6.0)
    ap.add_argument("--cfl", type=float, default=0.5)
    ap.add_argument("--limiter", default="mc", choices=["none", "minmod", "mc"])
    ap.add_argument("--report-every", type=int, default=50)
    ap.add_argument("--restart", default=None, help="write the final state to this .npz file")
    ap.add_argument("--resume", default=None,
                    help="start from this restart file, or from the last record of a history directory (case, N, dtype are taken from it)")
    ap.add_argument("--history", default=None, help="append records to this zarr v2 directory store (cubesphere.history)")
    ap.add_argument("--history-every", type=int, default=100, help="steps between history records")
    ap.add_argument("--second-order", action="store_true",
                    help="ghost values interpolated to our own extended grid (second order at the seams; mass drifts at truncation level)")
    ap.add_argument("--demo", default=None, choices=["diffusion"],
                    help="diffusion: heat from a checkerboard source on the top panel spreading over the seams (deck slides 12, 17)")
    ap.add_argument("--steps", type=int, default=1000, help="--demo diffusion: number of explicit steps")
    args = ap.parse_args(argv)

    import torch
    from pathlib import Path
    from . import api, cases, geometry, history, io

    if args.demo == "diffusion":
        return diffusion_demo(args)

    g = 2
    resumed, step0, time0 = None, 0, 0.0
    if args.resume:
        if Path(args.resume).is_dir():
            resumed, meta = history.ZarrHistory.open(args.resume).read(-1)
        else:
            resumed, meta = io.load_state(args.resume)
        if meta["tiles_per_edge"] != 1 or meta["halo"] != g or not meta.get("case"):
            raise ValueError(f"{args.resume}: need a whole-face state with halo {g} and a named case")
        args.case, args.N = meta["case"], meta["N"]
        args.dtype = "f32" if meta["dtype"] == "float32" else "f64"
        step0, time0 = meta["step"], meta["time"]
    case = cases.make_case(args.case, args.N)
    dt = cases.stable_dt(case, args.cfl)
    nsteps = max(1, int(round(args.hours * 3600.0 / dt)))
    ctx = api.Context(args.N, halo=g, dtype=args.dtype, limiter=args.limiter, advect_only=case.advect_only, axis=case.axis)
    ndt = np.float32 if args.dtype == "f32" else np.float64

    def padded(x):
        out = np.zeros((x.shape[0], args.N + 2 * g, args.N + 2 * g), dtype=ndt)
        out[:, g:-g, g:-g] = x
        return torch.from_numpy(out).cuda()

    if resumed is None:
        h, u, v = padded(case.h), padded(case.u1), padded(case.u2)
    else:
        h, u, v = (torch.from_numpy(np.ascontiguousarray(resumed[k], dtype=ndt)).cuda() for k in ("h", "u", "v"))
    if args.second_order:
        ctx.set_halo_interpolation(geometry.seam_interp_tables(args.N, g))
    if case.hs is not None:
        ctx.set_topography(padded(case.hs))
    ctx.halo_fill_state(h, u, v)
    d0 = ctx.diagnostics(h, u, v).cpu().numpy()
    h_start = h.clone() if args.case == "tc2" else None      # case 2 is steady: the initial state is the exact solution
    print(f"{args.case} C{args.N} {args.dtype}: dt = {dt:.2f} s, {nsteps} steps from step {step0}, "
          f"mass0 = {d0[0]:.6e}, energy0 = {d0[1]:.6e}")
    hist = None
    if args.history:
        hp = Path(args.history)
        hist = (history.ZarrHistory.open(hp) if (hp / ".zgroup").exists() else
                history.ZarrHistory.create(hp, N=args.N, halo=g, dtype=ndt, case=args.case))
        if (hist.N, hist.p, hist.g, hist.dtype) != (args.N, 1, g, np.dtype(ndt)):
            raise ValueError(f"{hp}: the store holds another grid or precision")
        if len(hist) == 0 or hist.read(-1)[1]["step"] < step0:
            hist.append({"h": h, "u": u, "v": v}, step=step0, time=time0, mass=d0[0], energy=d0[1])
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for k in range(1, nsteps + 1):
        ctx.step(h, u, v, dt)
        if hist is not None and (k % args.history_every == 0 or k == nsteps):
            d = ctx.diagnostics(h, u, v).cpu().numpy()
            hist.append({"h": h, "u": u, "v": v}, step=step0 + k, time=time0 + k * dt, mass=d[0], energy=d[1])
        if k % args.report_every == 0 or k == nsteps:
            d = ctx.diagnostics(h, u, v).cpu().numpy()
            print(f"  step {step0 + k:6d}  t = {(time0 + k * dt) / 3600.0:8.2f} h  mass drift = {(d[0] - d0[0]) / d0[0]:+.2e}  "
                  f"energy drift = {(d[1] - d0[1]) / d0[1]:+.2e}  h in [{float(h.min()):.2f}, {float(h.max()):.2f}]")
    torch.cuda.synchronize()
    sec = time.perf_counter() - t0
    print(f"{6.0 * args.N ** 2 * nsteps / sec / 1e9:.2f} G cell-updates/s including the diagnostics read-backs")
    if h_start is not None:
        c = slice(g, g + args.N)
        err = float(torch.linalg.norm((h - h_start)[:, c, c].double()) / torch.linalg.norm(h_start[:, c, c].double()))
        print(f"steady case: relative L2 error of h against the exact solution = {err:.3e}")
    if hist is not None:
        print(f"{len(hist)} records in {args.history}")
    if args.restart:
        io.save_state(args.restart, {"h": h, "u": u, "v": v}, N=args.N, halo=g, step=step0 + nsteps,
                      time=time0 + nsteps * dt, case=args.case)
        print(f"state written to {args.restart}")
    ctx.close()


def diffusion_demo(args):
    """Checkerboard heat source on face 0 for the first half of the run, then free diffusion: prints the heat on every
    face (the four neighbours of the top panel must receive equal shares) and the total, which the second half conserves."""
    import torch
    from . import api, geometry
    g, N = 2, args.N
    ctx = api.Context(N, halo=g, dtype=args.dtype)
    ndt = np.float32 if args.dtype == "f32" else np.float64
    t = geometry.centre_tangents(N)
    X, Y = np.meshgrid(t, t, indexing="xy")
    met = geometry.metric_terms(X, Y)
    kdt = 0.2 * (0.5 * np.pi / N) ** 2 / max(met["gi11"].max(), met["gi22"].max())
    src = np.zeros((6, N + 2 * g, N + 2 * g), dtype=ndt)
    k = (np.arange(N) * 4) // N
    src[0, g:g + N, g:g + N] = ((k[:, None] + k[None, :]) % 2) == 0
    dsrc = torch.from_numpy(src).cuda()
    a, b = ctx.new_field(), ctx.new_field()
    sg = torch.from_numpy(met["sg"]).cuda()

    def heat(T):
        return (T[:, g:g + N, g:g + N].double() * sg).sum(dim=(1, 2)).cpu().numpy()

    half = args.steps // 2
    for it in range(1, args.steps + 1):
        ctx.scalar_diffuse(a, b, kdt, source=dsrc if it <= half else None, dt_source=1.0)
        a, b = b, a
        if it % max(1, args.steps // 10) == 0:
            q = heat(a)
            print(f"  step {it:6d}  heat per face (share of total): " + " ".join(f"{x / q.sum():.4f}" for x in q) + f"   total {q.sum():.6e}")
    ctx.close()


if __name__ == "__main__":
    main()
