#!/usr/bin/env python
"""Benchmark of the hot path: explicit time steps of the cubed-sphere shallow-water solver.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

A "step" is one full SSP-RK3 time step (3 fused stage launches, halo fill, all three variables)
of the Galewsky barotropic-instability case at C1536 in fp32 -- the configuration
BASELINE.json's target is quoted on -- with synthetic (analytic) initial data.  The metric is
cell-updates/s = 6 N^2 x steps / seconds.  With N > 1 GPUs (launched by torchrun, one rank per
GPU) the sphere grows with the GPU count (weak scaling: C1536 / C2176 / C3072 / C4352 for
1 / 2 / 4 / 8 GPUs, 6 p^2 tiles block-partitioned over the ranks) and halos cross NVLink.

One JSON line is printed by rank 0; see DESIGN.md section 6 for what each key means.
`--impl reference` times the CPU oracle (the clean-room restatement; the upstream repository has
no runnable solver) with all host threads on the same workload, one step per "step".
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
PKG = ROOT / "sharding-the-sphere-fall-2025-jax-devlab-examples_b200"
for p in (str(PKG), str(ROOT)):
    if p not in sys.path:
        sys.path.insert(0, p)

import numpy as np  # noqa: E402

WEAK_N = {1: 1536, 2: 2176, 4: 3072, 8: 4352}
WEAK_P = {1: 1, 2: 1, 4: 2, 8: 2}
METRIC = "cell_updates_per_s"
UNIT = "cell-updates/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--N", type=int, default=0, help="cells per face edge; fixes the sphere (strong scaling) instead of the weak-scaling table")
    ap.add_argument("--dtype", default="f32", choices=["f32", "f64"])
    ap.add_argument("--case", default="galewsky")
    ap.add_argument("--kernel", default="auto")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-secondary", action="store_true", help="skip the fp64 side measurement")
    ap.add_argument("--new-state-sums", action="store_true",
                    help="per-step mass / energy of the NEW state (summed in the last stage) instead of the entry state (first stage)")
    ap.add_argument("--overlap", default="side", choices=["side", "rim_first", "none"],
                    help="N > 1: how a stage is split (rim kernel first on the whole GPU, or on a side stream beside the inner kernel)")
    ap.add_argument("--second-order", action="store_true",
                    help="interpolated (second-order) ghosts instead of the conservative copy (DESIGN.md 3.5); not the headline configuration")
    ap.add_argument("--tile-map", default="locality", choices=["locality", "block"],
                    help="N > 1: order of the tiles along the sharded axis (compact blocks, or the reference's plain order)")
    ap.add_argument("--halo-path", default="peer", choices=["peer", "nccl"],
                    help="N > 1: cross-GPU ghosts by one-sided peer stores (CUDA IPC over NVLink) or by pack / NCCL all-to-all / unpack")
    ap.add_argument("--no-parity", action="store_true", help="N > 1: skip the multi-GPU parity run that precedes the timing")
    return ap.parse_args()


def measured_peak_gbs():
    f = ROOT / "MEASURED_PEAKS.json"
    if f.exists():
        try:
            return float(json.loads(f.read_text())["hbm_gbs"]), "measured"
        except Exception:
            pass
    return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi SM clock / throttle-reason samples (one streaming process, 50 ms period) taken while
    the GPU measurements run: the timed steps, the per-launch stage timing and the end-to-end loop."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index, self.samples, self._proc, self._t = index, [], None, None

    def _run(self):
        for line in self._proc.stdout:
            parts = [x.strip() for x in line.split(",")]
            if len(parts) >= 6:
                self.samples.append(parts)

    def start(self):
        try:
            self._proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                           "-i", str(self.index), "-lms", "50"], stdout=subprocess.PIPE, text=True)
            self._t = threading.Thread(target=self._run, daemon=True)
            self._t.start()
        except Exception:
            self._proc = None
        return self

    def stop(self):
        if self._proc is not None:
            self._proc.terminate()
            try:
                self._proc.wait(timeout=5)
            except Exception:
                self._proc.kill()
            self._t.join(timeout=5)
            self._proc = None

    def __enter__(self):
        return self.start()

    def __exit__(self, *a):
        self.stop()

    def summary(self):
        mhz, mx, reasons = [], 0.0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for s in self.samples:
            try:
                mhz.append(float(s[0]))
                mx = max(mx, float(s[1]))
                for nm, val in zip(names, s[2:6]):
                    if val.lower().startswith("active"):
                        reasons.add(nm)
            except Exception:
                continue
        return {"sm_mhz": float(np.median(mhz)) if mhz else None, "sm_max_mhz": mx or None,
                "reasons": sorted(reasons), "samples": len(mhz)}


def initial_state(case_name, N):
    """Interior fields [6, N, N] fp64 of the named case, plus the Case object."""
    from cubesphere import cases
    case = cases.make_case(case_name, N)
    return case, cases.stable_dt(case, 0.5)


def faces_to_padded_tiles(field, p, g, dtype):
    """[6, N, N] interior -> [6 p^2, n + 2g, n + 2g] tiles with zero ghosts."""
    N = field.shape[-1]
    n = N // p
    out = np.zeros((6 * p * p, n + 2 * g, n + 2 * g), dtype=dtype)
    for f in range(6):
        for ty in range(p):
            for tx in range(p):
                out[f * p * p + ty * p + tx, g:g + n, g:g + n] = field[f, ty * n:(ty + 1) * n, tx * n:(tx + 1) * n]
    return out


# ----------------------------------------------------------------------------
# CPU arm (oracle; bench.py's cpu_baseline leg and the --impl reference arm)
# ----------------------------------------------------------------------------

def cpu_run(case, dt, steps, warmup, dtype="f64"):
    """Time `steps` steps of the C/OpenMP oracle after `warmup` steps, in the named precision (the fp32 build is the
    same source with the state, the tables and the per-cell arithmetic in float).
    Returns (cell-updates/s, threads, seconds)."""
    from oracle import cs_oracle_c as occ
    from oracle import cs_oracle_np as orc
    N, g = case.N, 2
    nd = np.float32 if dtype == "f32" else np.float64
    # every core this process may run on, whatever OMP_NUM_THREADS says (torchrun sets it to 1 for its workers)
    occ.set_threads(len(os.sched_getaffinity(0)), nd)
    hs = None
    if case.hs is not None:
        hs = orc.halo_fill_scalar(orc.pad(case.hs, g), g).astype(nd)
    co = occ.COracle(N, g, axis=case.axis, advect_only=case.advect_only, hs=hs, dtype=nd)
    st = [np.ascontiguousarray(orc.pad(x, g).astype(nd)) for x in (case.h, case.u1, case.u2)]
    occ.halo_fill(N, g, st[0], st[1], st[2])
    if warmup:
        co.step(*st, dt, warmup)
    t0 = time.perf_counter()
    co.step(*st, dt, steps)
    sec = time.perf_counter() - t0
    return 6.0 * N * N * steps / sec, co.threads, sec


def run_reference(args):
    """The CPU arm: the oracle's C + OpenMP port (the clean-room CPU implementation of the path; the upstream
    repository has no runnable solver) on all host threads.  Rank 0 only.  A step is one full time step; for
    multi-GPU workloads the sphere is capped at the 1-GPU size (C1536) so the run stays within minutes --
    cell-updates/s does not depend on the size once the state exceeds the CPU caches."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    N_job = args.N or WEAK_N.get(args.gpus, 1536)
    N = min(N_job, 1536)
    case, dt = initial_state(args.case, N)
    val, threads, sec = cpu_run(case, dt, args.steps, args.warmup, args.dtype)
    sample = f"{args.steps} full {args.dtype} steps of {args.case} C{N} after {args.warmup} warm-up, C + OpenMP oracle"
    if N != N_job:
        sample += f" (bounded sample of the C{N_job} job)"
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * sec / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": args.dtype, "data": "synthetic",
        "config": {"workload": f"{args.case} C{N_job}, SSP-RK3 (3 fused stages + halo) per step, {args.dtype}",
                   "N": N_job, "halo": 2, "run_as": f"C{N} {args.dtype} on the host cores"},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    _emit(line)


# ----------------------------------------------------------------------------
# GPU arm
# ----------------------------------------------------------------------------

def multi_gpu_parity(rank, world, local):
    """BASELINE configs[2] on the GPUs the job runs on: Williamson case 5 (flow over the mountain), C384, fp64,
    2 x 2 tiles per face block-partitioned over the ranks, 20 steps through the sharded stepper exactly as the
    timed loop drives it (rim / inner split, NCCL exchange overlapped, the step replayed as a CUDA graph), gathered
    to rank 0 and compared there with the C oracle on the host.  Also one sharded ghost fill of random numbers
    against the oracle's fill, bit for bit.  Returns the record on rank 0 (None elsewhere); the caller fails the
    run when a tolerance of BASELINE.json (rel-L2(h) <= 1e-12, mass 1e-13, halo bit-exact) is missed."""
    import torch
    import torch.distributed as dist
    from cubesphere import cases, sharded
    N, p, g, steps = 384, 2, 2, 20
    case = cases.make_case("tc5", N)
    dt = cases.stable_dt(case, 0.5)
    tang = None
    if rank == 0:
        from oracle import cs_oracle_np as orc            # the checker (rank 0 only)
        tang = orc.centre_tangents(N)
    box = [tang]
    dist.broadcast_object_list(box, src=0)                # every rank builds its tables from the oracle's tangents
    st = sharded.ShardedStepper(N, tiles_per_edge=p, halo=g, dtype="f64", axis=case.axis, device=local, rank=rank,
                                world=world, tangents=box[0])
    b, c = st.tile_begin, st.tile_count
    n = N // p

    def shard(face_field):                                # [6, N, N] interior -> this rank's padded tiles, ghosts zero
        out = np.zeros((c, n + 2 * g, n + 2 * g))
        for k, t in enumerate(st.tiles):
            f, ty, tx = t // (p * p), (t % (p * p)) // p, t % p
            out[k, g:g + n, g:g + n] = face_field[f, ty * n:(ty + 1) * n, tx * n:(tx + 1) * n]
        return out

    def gather(tensors):                                  # list of [c, m, m] device tensors -> [24, m, m] arrays on rank 0
        res = []
        for x in tensors:
            parts = [torch.empty_like(x) for _ in range(world)] if rank == 0 else None
            dist.gather(x, parts, dst=0)
            if rank == 0:
                by_pos = torch.cat(parts).cpu().numpy()
                by_tile = np.empty_like(by_pos)
                by_tile[st.tile_order] = by_pos              # position k of axis 0 holds tile tile_order[k]
                res.append(by_tile)
            else:
                res.append(None)
        return res

    # (1) ghost fill of random numbers, bit for bit
    rng = np.random.default_rng(20251118)
    rnd = [rng.standard_normal((6, N, N)) for _ in range(3)]
    dev = [torch.from_numpy(shard(x)).cuda() for x in rnd]
    st.halo_fill(*dev)
    got = gather(dev)
    halo_ok = None
    if rank == 0:
        want = orc.halo_fill_state(*(orc.pad(x, g) for x in rnd), g)
        halo_ok = all(np.array_equal(a, orc.faces_to_tiles(w, p, g)) for a, w in zip(got, want))
    # (2) 20 steps of case 5
    dev = [torch.from_numpy(shard(x)).cuda() for x in (case.h, case.u1, case.u2)]
    st.set_topography(torch.from_numpy(shard(case.hs)).cuda())
    st.halo_fill(*dev)
    m0 = st.diagnostics(*dev).cpu().numpy()
    for _ in range(steps):
        st.step(*dev, dt)
    m1 = st.diagnostics(*dev).cpu().numpy()
    got = gather(dev)
    rec = None
    if rank == 0:
        from oracle import cs_oracle_c as occ
        occ.set_threads(len(os.sched_getaffinity(0)))
        hs = orc.halo_fill_scalar(orc.pad(case.hs, g), g)
        co = occ.COracle(N, g, axis=case.axis, hs=hs)
        ref = [np.ascontiguousarray(orc.pad(x, g)) for x in (case.h, case.u1, case.u2)]
        occ.halo_fill(N, g, *ref)
        mo0 = co.diagnostics(*ref)
        co.step(*ref, dt, steps)
        mo1 = co.diagnostics(*ref)
        cc = slice(g, g + N)
        errs = []
        for a, w in zip(got, ref):
            a = orc.tiles_to_faces(a, p, g)
            errs.append(float(np.linalg.norm(a[:, cc, cc] - w[:, cc, cc]) / np.linalg.norm(w[:, cc, cc])))
        rec = {"config": f"tc5 C{N} fp64, {6 * p * p} tiles over {world} GPUs, {steps} steps, graph replay + overlap, vs oracle/cs_oracle.c",
               "rel_l2_h": errs[0], "rel_l2_u": errs[1], "rel_l2_v": errs[2],
               "mass_drift": float((m1[0] - m0[0]) / m0[0]), "mass_vs_oracle": float(abs(m1[0] - mo1[0]) / mo1[0]),
               "energy_vs_oracle": float(abs(m1[1] - mo1[1]) / mo1[1]), "oracle_mass_drift": float((mo1[0] - mo0[0]) / mo0[0]),
               "halo_bit_exact": bool(halo_ok)}
        rec["ok"] = bool(halo_ok and errs[0] <= 1e-12 and abs(rec["mass_drift"]) <= 1e-13 and rec["mass_vs_oracle"] <= 1e-13)
    st.close()
    del dev
    torch.cuda.empty_cache()
    return rec


def run_ours(args):
    import torch
    import torch.distributed as dist
    from cubesphere import api, partition, sharded

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE = {world}")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the hot path has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    # ---- multi-GPU parity on the hardware the numbers come from (before anything is timed) ------------------
    parity = None
    if world > 1 and not args.no_parity:
        parity = multi_gpu_parity(rank, world, local)
        flag = torch.tensor([1 if (rank != 0 or parity["ok"]) else 0], device="cuda")
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if int(flag.item()) != 1:
            if rank == 0:
                sys.stderr.write("multi-GPU parity FAILED: " + json.dumps(parity) + "\n")
                _emit({"metric": METRIC, "value": None, "n_gpus": world, "parity": parity, "error": "multi-GPU parity failed"})
            dist.destroy_process_group()
            os._exit(3)

    N = args.N or WEAK_N.get(world, 1536)
    p = WEAK_P.get(world, 1) if not args.N else (2 if world > 2 else 1)
    g = 2
    ndtype = np.float32 if args.dtype == "f32" else np.float64
    esize = 4 if args.dtype == "f32" else 8
    ranges = partition.block_ranges(6 * p * p, world)
    t0, tc = ranges[rank]
    # every rank builds only its own tiles of the initial condition (interior cells [tc, n, n])
    from cubesphere import cases
    order = partition.locality_order(p, world) if (args.tile_map == "locality" and world > 1) else np.arange(6 * p * p)
    case = cases.make_case(args.case, N, p=p, tiles=[int(t) for t in order[t0:t0 + tc]])
    dt = cases.stable_dt(case, 0.5)
    if world > 1:
        tdt = torch.tensor([dt], device="cuda", dtype=torch.float64)
        dist.all_reduce(tdt, op=dist.ReduceOp.MIN)
        dt = float(tdt.item())
    # conservation sums every step, taken inside the first stage's launches (the state each step starts from)
    stepper = sharded.ShardedStepper(N, tiles_per_edge=p, halo=g, dtype=args.dtype, axis=case.axis,
                                     advect_only=case.advect_only, device=local, rank=rank, world=world,
                                     kernel=args.kernel, sums_of_entry_state=not args.new_state_sums, halo_path=args.halo_path,
                                     overlap=False if args.overlap == "none" else args.overlap, tile_map=args.tile_map,
                                     second_order=args.second_order)
    assert (stepper.tile_begin, stepper.tile_count) == (t0, tc)

    def padded(field):
        out = np.zeros((field.shape[0], field.shape[1] + 2 * g, field.shape[2] + 2 * g), dtype=ndtype)
        out[:, g:-g, g:-g] = field
        return out

    host = [padded(x) for x in (case.h, case.u1, case.u2)]
    pinned = [torch.from_numpy(x).pin_memory() for x in host]
    dev = [x.cuda(non_blocking=True) for x in pinned]
    if case.hs is not None:
        stepper.set_topography(torch.from_numpy(padded(case.hs)).cuda())
    stepper.halo_fill(*dev)
    cells_total = 6.0 * N * N
    cells_local = float(tc) * stepper.n * stepper.n
    diag = torch.zeros(2, dtype=torch.float64, device="cuda")

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def one_step(with_sums=True):
        # N = 1: cs_step(..., diag_out) -- halo fills, three fused stages, conservation sums out of the last one
        if world == 1:
            stepper.ctx.step(*dev, dt, diag=diag if with_sums else None)
        else:
            # N > 1: every step's rank-local sums go to an on-device record; one all-reduce per timed region
            # (stepper.read_sums, inside the timing) turns the record into whole-sphere mass / energy per step
            stepper.step(*dev, dt, diag="history" if with_sums else None)

    def timed_steps(k, with_sums=True):
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        l0 = stepper.launches
        ev0.record()
        for _ in range(k):
            one_step(with_sums)
        if world > 1 and with_sums:
            timed_steps.record = stepper.read_sums_async()        # the one all-reduce of the region, inside the timing
        ev1.record()
        barrier()
        t_ms = ev0.elapsed_time(ev1)
        if world > 1:
            t = torch.tensor([t_ms], device="cuda", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            t_ms = float(t.item())
        return t_ms, stepper.launches - l0

    # ---- device-resident throughput: the whole step, conservation sums included ----------------------------
    for _ in range(args.warmup):
        one_step()
    barrier()
    clk = ClockSampler(local).start()
    time.sleep(0.15)                               # let the sampler come up; the GPU is warm from the warm-up steps
    for _ in range(args.warmup):
        one_step()
    ms, launches = timed_steps(args.steps)
    value = cells_total * args.steps / (ms * 1e-3)
    d_after = diag.cpu().numpy().tolist() if world == 1 else timed_steps.record[0][int(timed_steps.record[1].item()) - 1].cpu().numpy().tolist()
    finite = bool(torch.isfinite(dev[0]).all().item()) and bool(np.isfinite(d_after).all())
    for _ in range(4):                             # a new argument set: eager calls, then the graph capture, then a replay
        one_step(False)
    ms_nosums, _ = timed_steps(args.steps, with_sums=False)

    def peer_wait_timeouts():
        """One-sided exchange only: pulls (summed over the ranks) that gave up waiting for a peer so far.  Non-zero means
        ghosts went stale at some fill and every number measured after it is void."""
        if world == 1:
            return 0
        t = torch.tensor([stepper.peer_timeouts()], device="cuda", dtype=torch.int64)
        dist.all_reduce(t)
        return int(t.item())

    if peer_wait_timeouts():                       # the same count on every rank: all of them stop here
        raise RuntimeError("a one-sided halo wait timed out during the timed steps: the headline number is void")

    # ---- roofline of the dominant kernel (fused RK stage), timed per launch inside a CUDA graph --------------
    peak, peak_kind = measured_peak_gbs()
    roof = stepper.time_stage_kernel(dev, dt, reps=10)
    words = {True: 6, False: 9}                    # stage 1 reads qs, writes out; later stages also read q0
    alg_bytes = sum(words[a == 0.0] for a in roof["a"]) * esize * cells_local
    # the launches of the TIMED step: the stage that also forms the conservation sums is timed with them (at N > 1 the
    # stages are split into inner + rim launches and the whole-cell launch is timed without sums)
    stage_ms_nosums = [float(x) for x in roof["ms"]]
    if roof["fused_sums_us"] is not None:
        if args.new_state_sums:
            roof["ms"][-1] += roof["fused_sums_us"]["of_new_state_in_last_stage"] * 1e-3
        else:
            roof["ms"][0] += roof["fused_sums_us"]["of_entry_state_in_first_stage"] * 1e-3
    achieved = alg_bytes / (sum(roof["ms"]) * 1e-3) / 1e9
    # traffic: dram__bytes_read + dram__bytes_write per launch from the committed `ncu --set full` capture of
    # this kernel (profiles/r0N_traffic.json); only quoted when the capture's configuration is the one running
    traffic = None
    for name in ("r02_traffic.json", "r01_traffic.json"):
        try:
            tr = json.loads((ROOT / "profiles" / name).read_text())
            key = f"{args.dtype}_C{N}_tiles{tc}"
            if key in tr:
                traffic = tr[key]["dram_bytes_per_launch"]
                break
        except Exception:
            pass
    step_bytes = 24 * esize * cells_total / world
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic, "peak_kind": peak_kind, "kernel": roof["kernel"], "timing": roof["how"],
                "algorithmic_bytes_per_launch": alg_bytes / len(roof["ms"]),
                "algorithmic_bytes_per_cell_step": 24 * esize,
                "stage_ms": [float(x) for x in roof["ms"]],
                "stage_frac": [words[a == 0.0] * esize * cells_local / (t * 1e-3) / 1e9 / peak for a, t in zip(roof["a"], roof["ms"])],
                "stage_ms_without_sums": stage_ms_nosums,
                "frac_without_sums": alg_bytes / (sum(stage_ms_nosums) * 1e-3) / 1e9 / peak,
                "fused_sums_us": roof["fused_sums_us"],
                "frac_of_nominal_8TBs": achieved / 8000.0,
                "whole_step_frac": step_bytes / (ms / args.steps * 1e-3) / 1e9 / peak,
                "whole_step_frac_without_sums": step_bytes / (ms_nosums / args.steps * 1e-3) / 1e9 / peak}

    # ---- halo: the ghost-fill launch alone, and what the step pays beyond its three stage kernels -------
    halo = None
    if world == 1:
        # 50 fills captured in one CUDA graph: the replay time is device time, not host launch rate
        hev0, hev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for _ in range(3):
            stepper.ctx.halo_fill_state(*dev)
        torch.cuda.synchronize()
        hgraph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(hgraph):
            for _ in range(50):
                stepper.ctx.halo_fill_state(*dev)
        hgraph.replay()
        torch.cuda.synchronize()
        hev0.record()
        for _ in range(4):
            hgraph.replay()
        hev1.record()
        torch.cuda.synchronize()
        fill_us = hev0.elapsed_time(hev1) * 1e3 / 200
        del hgraph
        halo = {"fill_us": fill_us, "fills_per_step": len(roof["ms"]), "halo_us_per_step": fill_us * len(roof["ms"]),
                "exposed_us_per_step": max(0.0, (ms_nosums / args.steps - sum(roof["ms"])) * 1e3)}
    else:
        h = stepper.time_halo(dev, dt, reps=max(10, args.steps))
        halo = {"exchange_us": h["exchange_us"], "fills_per_step": len(roof["ms"]),
                "halo_us_per_step": h["exchange_us"] * len(roof["ms"]),
                "step_ms_without_exchange": h["step_ms_without_exchange"],
                "exposed_us_per_step": (ms_nosums / args.steps - h["step_ms_without_exchange"]) * 1e3,
                "what": "exchange_us = the cross-GPU part of one ghost fill (push + pull, or pack + all-to-all + unpack), alone on an idle GPU, max over ranks; "
                        "exposed = step time minus the time of the same step with the cross-GPU exchange left out",
                "path": stepper.halo_path}
        halo["peer_wait_timeouts"] = peer_wait_timeouts()      # 0 in a healthy run; otherwise this leg's numbers are void
        if "peer stores" in stepper.halo_path and not args.no_secondary and not args.second_order:
            # the comparison arm: the same job with pack -> NCCL all-to-all -> unpack
            alt = sharded.ShardedStepper(N, tiles_per_edge=p, halo=g, dtype=args.dtype, axis=case.axis,
                                         advect_only=case.advect_only, device=local, rank=rank, world=world,
                                         kernel=args.kernel, sums_of_entry_state=not args.new_state_sums, halo_path="nccl",
                                         tile_map=args.tile_map)
            adev = [x.clone() for x in dev]
            alt.halo_fill(*adev)
            for _ in range(5):
                alt.step(*adev, dt, diag="history")
            barrier()
            a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a0.record()
            for _ in range(args.steps):
                alt.step(*adev, dt, diag="history")
            alt.read_sums_async()
            a1.record()
            barrier()
            t = torch.tensor([a0.elapsed_time(a1) / args.steps], device="cuda", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ah = alt.time_halo(adev, dt, reps=max(10, args.steps))
            halo["nccl_arm"] = {"ms_per_step": float(t.item()), "exchange_us": ah["exchange_us"], "path": alt.halo_path}
            alt.close()
            del adev, alt
            torch.cuda.empty_cache()

    # ---- end to end through the host-buffer entry point: a ROUND TRIP (state in, new state + sums out) --------
    e2e = None
    if world == 1:
        ctx = stepper.ctx
        n_e2e = max(3, min(args.steps, 5))
        nbytes = sum(x.numel() * x.element_size() for x in pinned)
        ctx.step_host(*pinned, dt, nsteps=1, copy_back=True)             # warm-up (allocates the device copy)
        torch.cuda.synchronize()
        scratch = [x.clone().pin_memory() for x in pinned]               # the single-call legs advance a copy
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n_e2e):
            ctx.step_host(*scratch, dt, nsteps=1, copy_back=True)        # H2D state, step, D2H state + [mass, energy]
        e1.record()
        torch.cuda.synchronize()
        ms_single = e0.elapsed_time(e1) / n_e2e
        # The headline goes through the batch entry point: 8 members (2 distinct pinned states) stream through two
        # device slots, so the upload of member i + 1 and the download of member i - 1 overlap the step of member i.
        # Every member's state is copied host -> device, stepped, and its NEW STATE and [mass, energy] are copied
        # device -> host inside the timed region (step(state, dt) returns the state).
        ens = [scratch, [x.clone().pin_memory() for x in pinned]] * 4
        ctx.step_host_batch(ens, dt, nsteps=1, copy_back=True)           # warm-up (second slot, streams, graphs)
        t0_ = time.perf_counter()
        ctx.step_host_batch(ens, dt, nsteps=1, copy_back=True)
        ms_rt = (time.perf_counter() - t0_) * 1e3 / len(ens)
        t0_ = time.perf_counter()
        ctx.step_host_batch(ens, dt, nsteps=1, copy_back=False)
        ms_up = (time.perf_counter() - t0_) * 1e3 / len(ens)
        del ens, scratch
        e2e = {"value": cells_total / (ms_rt * 1e-3), "unit": UNIT, "h2d_bytes_per_step": nbytes,
               "d2h_bytes_per_step": nbytes + 16, "ms_per_step": ms_rt,
               "call": "cs_step_host_batch(copy_back=1): 8 members, 1 step each -- H2D state, halo fill, step, D2H new state "
                       "+ [mass, energy]; copies overlapped with the neighbours' steps; host wall clock around the blocking call",
               "sums_only": {"value": cells_total / (ms_up * 1e-3), "h2d_bytes_per_step": nbytes, "d2h_bytes_per_step": 16,
                             "ms_per_step": ms_up, "what": "same call with copy_back=0: only [mass, energy] come back"},
               "single_call": {"call": "cs_step_host(copy_back=1): one state per call, nothing overlapped; CUDA events",
                               "value": cells_total / (ms_single * 1e-3), "ms_per_step": ms_single}}
    else:
        e2e = stepper.time_e2e(pinned, dev, dt, max(3, min(args.steps, 5)), cells_total)
        e2e["peer_wait_timeouts"] = peer_wait_timeouts()       # ... and so are these, if any end-to-end fill gave up
        if e2e["peer_wait_timeouts"]:
            e2e["value"] = None

    clk.stop()

    # ---- fp64 side measurement (the metric is quoted for fp64 and fp32) ------------------------
    secondary = None
    if world == 1 and args.dtype == "f32" and not args.no_secondary:
        del dev
        torch.cuda.empty_cache()
        st64 = sharded.ShardedStepper(N, tiles_per_edge=p, halo=g, dtype="f64", axis=case.axis,
                                      advect_only=case.advect_only, device=local, rank=0, world=1, kernel=args.kernel,
                                      sums_of_entry_state=not args.new_state_sums)
        d64 = [torch.from_numpy(faces_to_padded_tiles(x, p, g, np.float64)).cuda() for x in (case.h, case.u1, case.u2)]
        st64.halo_fill(*d64)
        for _ in range(args.warmup):
            st64.ctx.step(*d64, dt, diag=diag)
        torch.cuda.synchronize()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record()
        for _ in range(args.steps):
            st64.ctx.step(*d64, dt, diag=diag)
        ev1.record()
        torch.cuda.synchronize()
        ms64 = ev0.elapsed_time(ev1)
        r64 = st64.time_stage_kernel(d64, dt, reps=10)
        if r64["fused_sums_us"] is not None:        # as above: the stage that forms the sums is timed with them
            k64, key64 = (-1, "of_new_state_in_last_stage") if args.new_state_sums else (0, "of_entry_state_in_first_stage")
            r64["ms"][k64] += r64["fused_sums_us"][key64] * 1e-3
        ach64 = sum(words[a == 0.0] for a in r64["a"]) * 8 * cells_local / (sum(r64["ms"]) * 1e-3) / 1e9
        secondary = {"dtype": "f64", "value": cells_total * args.steps / (ms64 * 1e-3), "unit": UNIT,
                     "ms_per_step": ms64 / args.steps, "roofline_frac": ach64 / peak, "achieved_GBps": ach64,
                     "stage_ms": [float(x) for x in r64["ms"]], "fused_sums_us": r64["fused_sums_us"],
                     "whole_step_frac": 24 * 8 * cells_total / (ms64 / args.steps * 1e-3) / 1e9 / peak}
        del d64, st64

    # ---- CPU baseline on rank 0 (bounded sample of the same workload, same precision) --------------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            # about 15 s of CPU work: time one probe step, then size the sample
            _, _, probe = cpu_run(case, dt, 1, 1, args.dtype)
            nsample = int(max(3, min(300, 15.0 / max(probe, 1e-3))))
            val, threads, sec = cpu_run(case, dt, nsample, 0, args.dtype)
            cpu = {"value": val, "unit": UNIT, "cores": threads, "kind": "port",
                   "sample": f"{nsample} full {args.dtype} steps of {args.case} C{N} after a warm-up, C + OpenMP oracle ({sec:.1f} s)"}
        except Exception as exc:  # pragma: no cover
            cpu = {"value": None, "unit": UNIT, "cores": 0, "kind": "port", "sample": f"failed: {exc}"}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong" if args.N else "weak", "vs_baseline": None,
            "dtype": args.dtype, "data": "synthetic",
            "config": {"workload": f"{args.case} C{N}, SSP-RK3 (3 fused stages + halo + mass/energy sums) per step, {args.dtype}",
                       "N": N, "tiles_per_edge": p, "halo": g, "tiles_per_gpu": tc, "parallelism": f"tiles/{world}",
                       "tile_map": stepper.tile_map + (f": {partition.cut_edges(stepper.tile_order, p, world)} of {12 * p * p} tile edges cross GPUs"
                                                       if world > 1 else ""),
                       "reductions": ("mass and energy every step: of the new state, summed inside the last stage launch"
                                      if args.new_state_sums else
                                      "mass and energy every step: of the state the step starts from, summed inside the first stage launch")
                                     + ("" if world == 1 else " (inner + rim kernels) into an on-device record; one all-reduce "
                                        "per timed region (inside the timing) combines the record across GPUs"),
                       "l2": "state (3 x %.0f MB x 3 buffers per GPU) is larger than L2; no flush needed" %
                             (cells_local * esize / 1e6),
                       "ghosts": "second order: interpolated + rotated, extended metric" if args.second_order else "copy + signed permutation (bit-exact, conservative)",
                       "finite": finite},
            "ms_per_step_without_sums": ms_nosums / args.steps, "diag": d_after,
            "roofline": roofline, "halo": halo, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches),
            "clocks": clk.summary(),
        }
        if parity is not None:
            line["parity"] = parity
        if secondary:
            line["fp64"] = secondary
        _emit(line)
    sys.stdout.flush()
    if world > 1:
        # a rank that is slow to tear NCCL down must not hold the job: hard exit after a grace period
        threading.Timer(60.0, lambda: os._exit(0)).start()
        stepper.close()
        dist.barrier()
        dist.destroy_process_group()
        os._exit(0)


_RESULT_FD = None


def _emit(line):
    """The one JSON line goes to the process's original stdout; everything else that writes to fd 1 during the run
    (NCCL prints its version banner there) has been sent to stderr."""
    data = (json.dumps(line) + "\n").encode()
    if _RESULT_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_RESULT_FD, data)


def main():
    global _RESULT_FD
    args = parse_args()
    sys.stdout.flush()
    _RESULT_FD = os.dup(1)
    os.dup2(2, 1)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
