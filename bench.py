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

def cpu_run(case, dt, steps, warmup):
    """Time `steps` steps of the C/OpenMP oracle after `warmup` steps.  Returns (cell-updates/s, threads, seconds)."""
    from oracle import cs_oracle_c as occ
    from oracle import cs_oracle_np as orc
    N, g = case.N, 2
    # every core this process may run on, whatever OMP_NUM_THREADS says (torchrun sets it to 1 for its workers)
    occ.set_threads(len(os.sched_getaffinity(0)))
    hs = None
    if case.hs is not None:
        hs = orc.halo_fill_scalar(orc.pad(case.hs, g), g)
    co = occ.COracle(N, g, axis=case.axis, advect_only=case.advect_only, hs=hs)
    st = [np.ascontiguousarray(orc.pad(x, g)) for x in (case.h, case.u1, case.u2)]
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
    val, threads, sec = cpu_run(case, dt, args.steps, args.warmup)
    sample = f"{args.steps} full fp64 steps of {args.case} C{N} after {args.warmup} warm-up"
    if N != N_job:
        sample += f" (bounded sample of the C{N_job} job)"
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * sec / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"{args.case} C{N_job}, SSP-RK3 (3 fused stages + halo) per step, {args.dtype}",
                   "N": N_job, "halo": 2, "run_as": f"C{N} fp64 on the host cores"},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    _emit(line)


# ----------------------------------------------------------------------------
# GPU arm
# ----------------------------------------------------------------------------

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

    N = args.N or WEAK_N.get(world, 1536)
    p = WEAK_P.get(world, 1) if not args.N else (2 if world > 2 else 1)
    g = 2
    tdtype = torch.float32 if args.dtype == "f32" else torch.float64
    ndtype = np.float32 if args.dtype == "f32" else np.float64
    esize = 4 if args.dtype == "f32" else 8
    ranges = partition.block_ranges(6 * p * p, world)
    t0, tc = ranges[rank]
    # every rank builds only its own tiles of the initial condition (interior cells [tc, n, n])
    from cubesphere import cases
    case = cases.make_case(args.case, N, p=p, tiles=range(t0, t0 + tc))
    dt = cases.stable_dt(case, 0.5)
    if world > 1:
        tdt = torch.tensor([dt], device="cuda", dtype=torch.float64)
        dist.all_reduce(tdt, op=dist.ReduceOp.MIN)
        dt = float(tdt.item())
    stepper = sharded.ShardedStepper(N, tiles_per_edge=p, halo=g, dtype=args.dtype, axis=case.axis,
                                     advect_only=case.advect_only, device=local, rank=rank, world=world,
                                     kernel=args.kernel)
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

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident throughput ------------------------------------------------------
    for _ in range(args.warmup):
        stepper.step(*dev, dt)
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    clk = ClockSampler(local).start()
    time.sleep(0.15)                               # let the sampler come up; the GPU is warm from the warm-up steps
    for _ in range(args.warmup):
        stepper.step(*dev, dt)
    barrier()
    l0 = stepper.launches
    ev0.record()
    for _ in range(args.steps):
        stepper.step(*dev, dt)
    ev1.record()
    barrier()
    ms = ev0.elapsed_time(ev1)
    launches = stepper.launches - l0
    if world > 1:
        t = torch.tensor([ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    value = cells_total * args.steps / (ms * 1e-3)
    finite = bool(torch.isfinite(dev[0]).all().item())

    # ---- roofline of the dominant kernel (fused RK stage), timed per launch ------------------
    peak, peak_kind = measured_peak_gbs()
    roof = stepper.time_stage_kernel(dev, dt, reps=max(3, min(args.steps, 10)))
    words = {True: 6, False: 9}                    # stage 1 reads qs, writes out; later stages also read q0
    alg_bytes = sum(words[a == 0.0] for a in roof["a"]) * esize * cells_local
    achieved = alg_bytes / (sum(roof["ms"]) * 1e-3) / 1e9
    # traffic: dram__bytes_read + dram__bytes_write per launch from the committed `ncu --set full` capture of
    # this kernel (profiles/r01_traffic.json); only quoted when the capture's configuration is the one running
    traffic = None
    try:
        tr = json.loads((ROOT / "profiles" / "r01_traffic.json").read_text())
        key = f"{args.dtype}_C{N}_tiles{tc}"
        if key in tr:
            traffic = tr[key]["dram_bytes_per_launch"]
    except Exception:
        pass
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic, "peak_kind": peak_kind, "kernel": roof["kernel"],
                "algorithmic_bytes_per_launch": alg_bytes / len(roof["ms"]),
                "algorithmic_bytes_per_cell_step": 24 * esize,
                "stage_ms": [float(x) for x in roof["ms"]], "frac_of_nominal_8TBs": achieved / 8000.0,
                "whole_step_frac": 24 * esize * cells_total / world / (ms / args.steps * 1e-3) / 1e9 / peak}

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
                "exposed_us_per_step": max(0.0, (ms / args.steps - sum(roof["ms"])) * 1e3)}

    # ---- end to end through the host-buffer entry point ---------------------------------------
    e2e = None
    if world == 1:
        ctx = stepper.ctx
        n_e2e = max(3, min(args.steps, 5))
        ctx.step_host(*pinned, dt, nsteps=1, copy_back=False)            # warm-up (allocates the device copy)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n_e2e):
            ctx.step_host(*pinned, dt, nsteps=1, copy_back=False)        # H2D state, step, D2H [mass, energy]
        e1.record()
        torch.cuda.synchronize()
        ms_a = e0.elapsed_time(e1) / n_e2e
        e0.record()
        for _ in range(n_e2e):
            ctx.step_host(*pinned, dt, nsteps=1, copy_back=True)         # ... and the whole state back
        e1.record()
        torch.cuda.synchronize()
        ms_b = e0.elapsed_time(e1) / n_e2e
        nbytes = sum(x.numel() * x.element_size() for x in pinned)
        # The headline goes through the batch entry point: 8 members (2 distinct pinned states) stream through two
        # device slots, so the upload of member i + 1 and the download of member i - 1 overlap the step of member i.
        # Every member's state is copied host -> device and its [mass, energy] (round_trip: its whole state)
        # device -> host inside the timed region.  The one-state-per-call number is reported beside it.
        ens = [pinned, [x.clone().pin_memory() for x in pinned]] * 4
        ctx.step_host_batch(ens, dt, nsteps=1, copy_back=False)          # warm-up (second slot, streams, graphs)
        t0 = time.perf_counter()
        ctx.step_host_batch(ens, dt, nsteps=1, copy_back=False)
        ms_c = (time.perf_counter() - t0) * 1e3 / len(ens)
        t0 = time.perf_counter()
        ctx.step_host_batch(ens, dt, nsteps=1, copy_back=True)
        ms_d = (time.perf_counter() - t0) * 1e3 / len(ens)
        del ens
        e2e = {"value": cells_total / (ms_c * 1e-3), "unit": UNIT, "h2d_bytes_per_step": nbytes,
               "d2h_bytes_per_step": 16, "ms_per_step": ms_c,
               "call": "cs_step_host_batch: 8 members, 1 step each (H2D state, halo fill, step, D2H [mass, energy]), "
                       "copies overlapped with the neighbours' steps; host wall clock around the blocking call",
               "round_trip": {"value": cells_total / (ms_d * 1e-3), "h2d_bytes_per_step": nbytes,
                              "d2h_bytes_per_step": nbytes + 16, "ms_per_step": ms_d},
               "single_call": {"call": "cs_step_host: one state per call, nothing overlapped; CUDA events",
                               "value": cells_total / (ms_a * 1e-3), "ms_per_step": ms_a,
                               "round_trip": {"value": cells_total / (ms_b * 1e-3), "ms_per_step": ms_b}}}
    else:
        e2e = stepper.time_e2e(pinned, dev, dt, max(3, min(args.steps, 5)), cells_total)

    clk.stop()

    # ---- fp64 side measurement (the metric is quoted for fp64 and fp32) ------------------------
    secondary = None
    if world == 1 and args.dtype == "f32" and not args.no_secondary:
        del dev
        torch.cuda.empty_cache()
        st64 = sharded.ShardedStepper(N, tiles_per_edge=p, halo=g, dtype="f64", axis=case.axis,
                                      advect_only=case.advect_only, device=local, rank=0, world=1, kernel=args.kernel)
        d64 = [torch.from_numpy(faces_to_padded_tiles(x, p, g, np.float64)).cuda() for x in (case.h, case.u1, case.u2)]
        st64.halo_fill(*d64)
        for _ in range(args.warmup):
            st64.step(*d64, dt)
        torch.cuda.synchronize()
        ev0.record()
        for _ in range(args.steps):
            st64.step(*d64, dt)
        ev1.record()
        torch.cuda.synchronize()
        ms64 = ev0.elapsed_time(ev1)
        r64 = st64.time_stage_kernel(d64, dt, reps=3)
        ach64 = sum(words[a == 0.0] for a in r64["a"]) * 8 * cells_local / (sum(r64["ms"]) * 1e-3) / 1e9
        secondary = {"dtype": "f64", "value": cells_total * args.steps / (ms64 * 1e-3), "unit": UNIT,
                     "ms_per_step": ms64 / args.steps, "roofline_frac": ach64 / peak, "achieved_GBps": ach64}
        del d64, st64

    # ---- CPU baseline on rank 0 (bounded sample of the same workload) --------------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            # about 15 s of CPU work: time one probe step, then size the sample
            _, _, probe = cpu_run(case, dt, 1, 1)
            nsample = int(max(3, min(300, 15.0 / max(probe, 1e-3))))
            val, threads, sec = cpu_run(case, dt, nsample, 0)
            cpu = {"value": val, "unit": UNIT, "cores": threads, "kind": "port",
                   "sample": f"{nsample} full fp64 steps of {args.case} C{N} after a warm-up, C + OpenMP oracle ({sec:.1f} s)"}
        except Exception as exc:  # pragma: no cover
            cpu = {"value": None, "unit": UNIT, "cores": 0, "kind": "port", "sample": f"failed: {exc}"}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong" if args.N else "weak", "vs_baseline": None,
            "dtype": args.dtype, "data": "synthetic",
            "config": {"workload": f"{args.case} C{N}, SSP-RK3 (3 fused stages + halo) per step, {args.dtype}",
                       "N": N, "tiles_per_edge": p, "halo": g, "tiles_per_gpu": tc, "parallelism": f"tiles/{world}",
                       "l2": "state (3 x %.0f MB x 3 buffers per GPU) is larger than L2; no flush needed" %
                             (cells_local * esize / 1e6),
                       "finite": finite},
            "roofline": roofline, "halo": halo, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches),
            "clocks": clk.summary(),
        }
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
