"""One-process-per-GPU driver of the step over a block partition of the tile axis.

world == 1: every tile lives on one device and the whole step is ``cs_step`` (halo included).
world  > 1: each rank owns ``6 p^2 / world`` consecutive POSITIONS of axis 0 (the reference's ``P('tiles')`` block
partition, JAX-DevLab-Examples.py:40-57, 77-79); which tile sits at which position is ``tile_map`` ("locality": compact
blocks found by ``partition.locality_order``; "block": tile k at position k).  Per RK stage

    side stream                                                   main stream
    rim cells (what ghosts depend on)   stage kernel, part 1      inner cells: stage kernel, part 0
    same-rank ghosts                    ghost-fill kernel         (almost all of the time; leaves the side
    cross-GPU ghosts, one-sided:        push kernel (stores into   stream a few SMs)
      the peers' receive buffers over NVLink + flag), pull kernel
      (waits for its flags, scatters)   -- or, ``halo_path="nccl"``: pack, NCCL all-to-all, unpack

Inner cells never read ghosts: stage k+1's inner kernel waits for stage k's inner and rim kernels only, the exchange has
a whole stage of slack; the streams join at the end of the step, and the whole step is captured into one CUDA graph.
Conservation sums come out of a stage's two launches (four rank-local doubles) and are combined across ranks either
every step (``diag=tensor``) or once per read of an on-device record (``diag="history"``, :meth:`read_sums`).
"""
from __future__ import annotations

import os

import torch
import torch.distributed as dist

from . import api, topology
from .partition import block_ranges, locality_order

RK_PAIRS = {
    "euler": ((0.0, 1.0),),
    "ssprk2": ((0.0, 1.0), (0.5, 0.5)),
    "ssprk3": ((0.0, 1.0), (0.75, 0.25), (1.0 - 2.0 / 3.0, 2.0 / 3.0)),
}


class ShardedStepper:
    def __init__(self, N, tiles_per_edge=1, halo=2, dtype="f32", rk="ssprk3", limiter="mc", advect_only=False,
                 axis=(0.0, 0.0, 1.0), device=0, rank=0, world=1, kernel="auto", tangents=None, overlap=True,
                 group=None, use_graph=True, sums_of_entry_state=False, halo_path="peer", tile_map="locality",
                 second_order=False):
        # second_order: ghost values interpolated to this panel's own extended grid instead of copied --
        # second order at the seams, mass no longer to round-off.  On several GPUs it needs the one-sided exchange
        # (raw cells travel, the receiver interpolates).
        self.second_order = bool(second_order)
        # sums_of_entry_state: step(..., diag) reports mass / energy of the state the step STARTS from, summed inside
        # the first stage's launches (a few instructions per cell); default: of the new state, inside the last stage's
        self.sums_of_entry_state = bool(sums_of_entry_state)
        if overlap is True:
            overlap = "side"        # measured at 2 GPUs, C2176 fp32: 0.314 ms per step against 0.340 for "rim_first"
        if overlap not in (False, "rim_first", "side"):
            raise ValueError("overlap must be False, True / 'rim_first' or 'side'")
        self.rank, self.world, self.rk, self.overlap, self.group = rank, world, rk, overlap, group
        # world > 1: a step is ~30 host calls (kernel launches, stream waits, one NCCL all-to-all per stage);
        # captured once into a CUDA graph per (dt, state tensors) and replayed, so the host never paces the GPUs
        self.use_graph = use_graph and world > 1
        self._graphs = {}
        self._eager_calls = {}
        self.ranges = block_ranges(topology.num_tiles(tiles_per_edge), world)
        self.tile_begin, self.tile_count = self.ranges[rank]          # positions along axis 0 (the P('tiles') blocks)
        # which tile sits at which position: "block" = the reference's order (tile k at position k), "locality" = an
        # order whose blocks are spatially compact (fewer tile edges, and fewer peers, across GPUs), or an explicit table
        if isinstance(tile_map, str):
            if tile_map not in ("block", "locality"):
                raise ValueError("tile_map must be 'block', 'locality' or a permutation of the tile ids")
            order = locality_order(tiles_per_edge, world) if tile_map == "locality" and world > 1 else None
        else:
            order = tile_map
        self.tile_map = "block" if order is None else "locality" if isinstance(tile_map, str) else "custom"
        self.ctx = api.Context(N, tiles_per_edge=tiles_per_edge, halo=halo, dtype=dtype, rk=rk, limiter=limiter,
                               advect_only=advect_only, axis=axis, device=device, kernel=kernel, tangents=tangents,
                               sums_of_entry_state=sums_of_entry_state, tile_order=order,
                               tile_begin=self.tile_begin, tile_count=self.tile_count if world > 1 else 0)
        self.n, self.m, self.device = self.ctx.n, self.ctx.m, device
        self.tiles = [int(t) for t in self.ctx.tiles]                   # tile ids of this rank's shard, in shard order
        self.tile_order = self.ctx.tile_order
        self._scratch = None
        self._extra_launches = 0
        self._sums_where = os.environ.get("CS_SUMS_WHERE", "own_stream")       # tuning knob: own_stream | comm_stream
        self._skip_exchange = False      # measurement only (time_halo): a step with the cross-GPU exchange left out
        if world == 1 and self.second_order:
            from . import geometry
            self.ctx.set_halo_interpolation(geometry.seam_interp_tables(N, halo, tiles_per_edge))
        if world > 1:
            if self.second_order:
                if halo_path != "peer":
                    raise ValueError("second_order on several GPUs needs halo_path='peer'")
                from .partition import InterpExchangePlan
                plan = InterpExchangePlan(N, tiles_per_edge, halo, rank, self.ranges, self.ctx.tile_order)
                self.ctx.halo_interp_sharded(plan)
                self.send_counts, self.recv_counts = plan.send_counts, plan.recv_counts
            else:
                self.send_counts, self.recv_counts = self.ctx.halo_plan([b for b, _ in self.ranges],
                                                                        [c for _, c in self.ranges])
            dev = f"cuda:{device}"
            self.sendbuf = torch.zeros(max(1, 3 * sum(self.send_counts)), dtype=self.ctx.dtype, device=dev)
            self.recvbuf = torch.zeros(max(1, 3 * sum(self.recv_counts)), dtype=self.ctx.dtype, device=dev)
            # CS_COMM_PRIORITY=1 gives the side stream the higher priority (its short rim kernel is then placed on the
            # SMs before the blocks of the inner kernel); measured: no gain (2 GPUs 0.299 vs 0.294 ms, 8 GPUs 0.309 vs 0.310)
            prio = -1 if os.environ.get("CS_COMM_PRIORITY", "0") != "0" else 0
            self.comm_stream = torch.cuda.Stream(device=device, priority=prio)
            self.sums_stream = torch.cuda.Stream(device=device)
            self.qa = [self.ctx.new_field() for _ in range(3)]
            self.qb = [self.ctx.new_field() for _ in range(3)]
            self._diag = torch.zeros(2, dtype=torch.float64, device=dev)
            self._sums4 = torch.zeros(4, dtype=torch.float64, device=dev)    # [inner mass, energy, rim mass, energy]
            # deferred mode (diag="history"): every step's rank-local sums are appended here on the device and ONE
            # all-reduce combines the whole record when it is read (read_sums)
            self.history_capacity = 4096
            self._hist = torch.zeros(self.history_capacity, 4, dtype=torch.float64, device=dev)
            self._hist_idx = torch.zeros(1, dtype=torch.int64, device=dev)
        self.halo_path = "single device: one ghost-fill launch"
        if world > 1:
            # cross-GPU ghosts: one-sided peer stores over NVLink (CUDA IPC mappings, one push and one pull kernel per
            # fill) or, as the comparison arm and the fallback, pack -> NCCL all-to-all -> unpack
            if halo_path not in ("peer", "nccl"):
                raise ValueError(f"halo_path must be 'peer' or 'nccl', got {halo_path!r}")
            self._peer = False
            if halo_path == "peer":
                self._peer = self._connect_peers()
                if self.second_order and not self._peer:
                    raise RuntimeError("second_order needs the one-sided exchange, and the peers' buffers could not be mapped")
            self.halo_path = ("peer stores into the neighbours' receive buffers (CUDA IPC over NVLink) + flag, then pull"
                              if self._peer else "pack -> NCCL all_to_all_single -> unpack")
            if self._peer and "CS_MARCH_RESERVE_SMS" not in os.environ:
                # the side stream runs four short kernels of ours and no NCCL kernel: a few SMs are enough (12 with
                # NCCL).  Measured, C2176 fp32 per GPU: 2 GPUs 0.296 ms per step with 4 SMs, 0.307 with 8; 8 GPUs (seven
                # peers to wait for, the exchange 33 us instead of 19) 0.306 with 8, 0.310 with 4.
                self.ctx.set_option("reserve_sms", 4 if world <= 4 else 8)

    # -- bookkeeping ------------------------------------------------------------------------
    @property
    def launches(self):
        return self.ctx.launches + self._extra_launches

    def peer_timeouts(self):
        """How many one-sided halo waits on this rank gave up (synchronises).  0 in a healthy run."""
        return self.ctx.halo_peer_timeouts() if getattr(self, "_peer", False) else 0

    def check_peers(self):
        """Raise if a pull of the one-sided exchange has given up waiting for a peer (option ``peer_timeout_ms``, default
        5 s): from that fill on the ghosts are stale and no later pull waits, so the results are not to be used."""
        n = self.peer_timeouts()
        if n:
            raise RuntimeError(f"rank {self.rank}: {n} one-sided halo wait(s) timed out -- a peer did not issue its ghost "
                               "fill (ranks out of step, or a rank died); results since then are invalid")

    def close(self):
        # graphs hold NCCL work: drop them (and drain the device) before the process group goes away
        self._graphs.clear()
        torch.cuda.synchronize(self.device)
        if self.peer_timeouts():
            import warnings
            warnings.warn(f"rank {self.rank}: one-sided halo waits timed out during this run (see check_peers)")
        if self.world > 1 and getattr(self, "_peer", False) and dist.is_initialized():
            dist.barrier(group=self.group)                 # peers hold mappings of this context's buffers
        self.ctx.close()

    def _connect_peers(self):
        """Exchange IPC handles and receive counts, map the peers' buffers.  False (-> NCCL path) when a rank cannot."""
        ok, handle = 1, b""
        try:
            handle = self.ctx.halo_peer_alloc()
        except Exception:
            ok = 0
        everyone = [None] * self.world
        dist.all_gather_object(everyone, (ok, handle, list(self.recv_counts)), group=self.group)
        if not all(e[0] for e in everyone):
            return False
        # my block inside rank r's receive buffer starts after the blocks of the ranks before me
        remote = [sum(everyone[r][2][:self.rank]) for r in range(self.world)]
        try:
            self.ctx.halo_peer_connect(self.rank, [e[1] for e in everyone], remote)
        except Exception:
            ok = 0
        flags = [None] * self.world
        dist.all_gather_object(flags, ok, group=self.group)
        return all(flags)

    # -- halo -------------------------------------------------------------------------------
    def _exchange(self, h, u, v, nfields=3):
        """pack -> all-to-all -> unpack on the current stream (remote ghosts only)."""
        if self._skip_exchange:
            return
        if self._peer:
            self.ctx.halo_push(h, u, v, nfields)
            self.ctx.halo_pull(h, u, v, nfields)
            return
        self.ctx.halo_pack(h, u, v, self.sendbuf, nfields)
        ns = nfields * sum(self.send_counts)
        nr = nfields * sum(self.recv_counts)
        dist.all_to_all_single(self.recvbuf[:nr], self.sendbuf[:ns],
                               output_split_sizes=[nfields * c for c in self.recv_counts],
                               input_split_sizes=[nfields * c for c in self.send_counts], group=self.group)
        self.ctx.halo_unpack(self.recvbuf, h, u, v, nfields)

    def halo_fill(self, h, u, v):
        self.ctx.halo_fill_state(h, u, v)
        if self.world > 1:
            self._exchange(h, u, v)
        return h, u, v

    def set_topography(self, hs):
        """``hs``: this rank's tiles of the bottom topography (interiors set; ghosts are filled here,
        from same-rank neighbours by the halo kernel and from peers by a one-field exchange)."""
        if self.world > 1:
            hs = hs.clone()
            self.ctx.halo_fill_scalar(hs)
            self._exchange(hs, None, None, nfields=1)
        self.ctx.set_topography(hs)

    # -- step -------------------------------------------------------------------------------
    def step(self, h, u, v, dt, diag=None):
        """One time step in place.  ``diag`` (device fp64 tensor [2], optional) receives mass and total energy, whole
        sphere: the sums come out of a stage's launches (no extra pass over the state), and across ranks they are
        combined by one all-reduce of four doubles per step.  ``diag="history"`` (world > 1) defers the collective:
        the rank-local sums of every step are appended to an on-device record and :meth:`read_sums` combines the whole
        record in one all-reduce."""
        if isinstance(diag, str) and diag != "history":
            raise ValueError("diag must be a device tensor, None or 'history'")
        if self.world == 1:
            if isinstance(diag, str):
                raise ValueError("diag='history' is the multi-GPU deferred mode; pass a tensor on one GPU")
            self.ctx.step(h, u, v, dt, diag=diag)
            return h, u, v
        if not self.use_graph:
            return self._step_sharded(h, u, v, dt, diag)
        key = (float(dt), h.data_ptr(), u.data_ptr(), v.data_ptr(),
               0 if diag is None else -1 if isinstance(diag, str) else diag.data_ptr())
        graph = self._graphs.get(key)
        if graph is None:
            # the first two calls run eagerly (NCCL warm-up, lazy initialisation), the third is captured
            n = self._eager_calls.get(key, 0)
            if n < 2:
                self._eager_calls[key] = n + 1
                return self._step_sharded(h, u, v, dt, diag)
            torch.cuda.synchronize(self.device)
            graph = torch.cuda.CUDAGraph()
            launches0 = self.ctx.launches
            with torch.cuda.graph(graph):
                self._step_sharded(h, u, v, dt, diag)
            self._graph_launches = self.ctx.launches - launches0     # kernels of ours inside one replay
            self._graphs[key] = graph
            self._extra_launches -= self._graph_launches              # counted once by the capture itself
            # capture does not execute: fall through and replay once so this call does advance the state
        graph.replay()
        self._extra_launches += self._graph_launches
        return h, u, v

    def _step_sharded(self, h, u, v, dt, diag=None):
        q0 = (h, u, v)
        cur = q0
        pairs = RK_PAIRS[self.rk]
        main = torch.cuda.current_stream(self.device)
        ev_rim = ev_inner = ev_exch = None
        pending_sums = joined_sums = False
        for k, (a, b) in enumerate(pairs):
            last = k == len(pairs) - 1 and len(pairs) > 1
            out = q0 if last else (self.qa if k % 2 == 0 else self.qb)
            # the last stage also sums mass / energy of its output -- or the first one those of its input
            sin = self.sums_of_entry_state
            sums = diag is not None and k == (0 if sin else len(pairs) - 1)
            if self.overlap == "rim_first":
                # main stream: the rim cells first, on the whole GPU (a few microseconds: they are 8 / n of the cells),
                # then the inner cells; side stream: same-rank ghosts and the cross-GPU exchange of the fresh rim cells,
                # under the cover of the inner kernel (which leaves it a few SMs).  Only the NEXT stage's rim kernel needs
                # what the exchange brings, so it has a whole stage of slack; the streams join at the end of the step.
                if ev_exch is not None:
                    main.wait_event(ev_exch)                                   # the rim reads remote ghosts of `cur`
                self.ctx.rk_stage(q0, cur, out, a, b, dt, part=1, diag=self._sums4[2:] if sums else None, sums_of_input=sin)
                ev_rim = torch.cuda.Event()
                ev_rim.record(main)
                with torch.cuda.stream(self.comm_stream):
                    self.comm_stream.wait_event(ev_rim)
                    self.ctx.halo_fill_state(*out)
                    self._exchange(*out)
                    ev_exch = torch.cuda.Event()
                    ev_exch.record(self.comm_stream)
                self.ctx.rk_stage(q0, cur, out, a, b, dt, part=0, diag=self._sums4[:2] if sums else None, sums_of_input=sin)
                if sums and sin and k < len(pairs) - 1:
                    # entry-state sums are complete: combine / record them on a stream of their own, under stages 2, 3
                    ev_inner = torch.cuda.Event()
                    ev_inner.record(main)
                    self.sums_stream.wait_event(ev_inner)
                    with torch.cuda.stream(self.sums_stream):
                        self._combine_sums(diag)
                    joined_sums = True
            elif self.overlap:
                # (overlap="side", the round-1 arrangement, kept for comparison)
                # side stream: rim cells -> same-rank ghosts -> exchange; main stream: the inner cells.
                # Inner cells never read a ghost cell, so stage k+1's inner kernel only waits for stage k's
                # inner and rim kernels -- not for its exchange, which gets a whole stage of slack and is
                # needed by stage k+1's rim kernel alone.  The two streams join at the end of the step.
                with torch.cuda.stream(self.comm_stream):
                    if ev_inner is None:
                        self.comm_stream.wait_stream(main)                     # whatever came before the step
                    else:
                        self.comm_stream.wait_event(ev_inner)                  # rim reads inner cells next to it
                        if pending_sums:
                            # entry-state sums came out of stage 1 (inner + rim kernels, both finished by now)
                            mode = self._sums_where
                            if mode == "own_stream":
                                self.sums_stream.wait_stream(self.comm_stream)
                                with torch.cuda.stream(self.sums_stream):
                                    self._combine_sums(diag)
                                joined_sums = True
                            elif mode == "comm_stream":
                                self._combine_sums(diag)
                            pending_sums = False
                    self.ctx.rk_stage(q0, cur, out, a, b, dt, part=1, diag=self._sums4[2:] if sums else None, sums_of_input=sin)
                    ev_rim_new = torch.cuda.Event()
                    ev_rim_new.record(self.comm_stream)
                    self.ctx.halo_fill_state(*out)
                    self._exchange(*out)
                if ev_rim is not None:
                    main.wait_event(ev_rim)                                    # inner reads the previous rim cells
                self.ctx.rk_stage(q0, cur, out, a, b, dt, part=0, diag=self._sums4[:2] if sums else None, sums_of_input=sin)
                ev_inner = torch.cuda.Event()
                ev_inner.record(main)
                ev_rim = ev_rim_new
                pending_sums = pending_sums or (sums and sin and k < len(pairs) - 1)
            else:
                if sums:
                    self._sums4[2:].zero_()
                self.ctx.rk_stage(q0, cur, out, a, b, dt, part=2, diag=self._sums4[:2] if sums else None, sums_of_input=sin)
                self.ctx.halo_fill_state(*out)
                self._exchange(*out)
            cur = out
        if self.overlap:
            main.wait_stream(self.comm_stream)
            if joined_sums:
                main.wait_stream(self.sums_stream)
        if diag is not None and not (self.overlap and self.sums_of_entry_state and len(pairs) > 1):
            self._combine_sums(diag)                       # sums of the new state: only now are they complete
        if len(pairs) == 1:
            for x, y in zip(q0, cur):
                x.copy_(y)
        return h, u, v

    def _combine_sums(self, diag):
        """Rank-local (inner, rim) sums -> whole sphere: one all-reduce of four doubles, then inner + rim.
        ``diag == "history"``: no collective now -- append the local sums to the on-device record instead."""
        if isinstance(diag, str):
            self._hist.index_copy_(0, self._hist_idx % self.history_capacity, self._sums4.unsqueeze(0))
            self._hist_idx.add_(1)
            return
        dist.all_reduce(self._sums4, group=self.group)
        torch.add(self._sums4[:2], self._sums4[2:], out=diag)

    def read_sums_async(self):
        """Deferred mode, no host synchronisation: enqueues ONE all-reduce over the whole on-device record and restarts
        the record; returns ``(sums [capacity, 2], count [1])`` device tensors -- rows ``[:count]`` are (mass, energy) of
        the whole sphere for every step taken with ``diag="history"`` since the last read."""
        rec = self._hist.clone()
        count = self._hist_idx.clone()
        self._hist_idx.zero_()
        if self.world > 1:
            dist.all_reduce(rec, group=self.group)
        return rec[:, :2] + rec[:, 2:], count

    def read_sums(self):
        """:meth:`read_sums_async`, then the rows that were recorded, as a ``[steps, 2]`` tensor (synchronises)."""
        sums, count = self.read_sums_async()
        n = int(count.item())
        self.check_peers()
        if n > self.history_capacity:
            raise RuntimeError(f"{n} steps recorded, the history holds {self.history_capacity}: read it more often")
        return sums[:n]

    def diagnostics(self, h, u, v):
        """[mass, energy] of the whole sphere on every rank (fp64)."""
        out = self.ctx.diagnostics(h, u, v)
        if self.world > 1:
            dist.all_reduce(out, group=self.group)
        return out

    def write_history(self, hist, h, u, v, *, step, time):
        """Append the current state to a :class:`cubesphere.history.ZarrHistory` (every rank opens the same store on a
        shared file system).  Each rank writes the chunk files of the tiles it owns straight from its shard -- the
        fields are never gathered; rank 0 adds the scalars (with the whole-sphere mass / energy) and publishes the record
        after a barrier.  Synchronises; not for the timed path."""
        from . import history
        sums = self.diagnostics(h, u, v).cpu().numpy()
        barrier = (lambda: dist.barrier(group=self.group)) if self.world > 1 else (lambda: None)
        return history.write_sharded(hist, {"h": h, "u": u, "v": v}, tile_ids=self.tiles, rank=self.rank, barrier=barrier,
                                     step=step, time=time, mass=sums[0], energy=sums[1])

    # -- measurement helpers used by bench.py ---------------------------------------------------
    def time_stage_kernel(self, state, dt, reps=10, rounds=3):
        """Device time of each fused-stage launch alone (all owned cells, no halo): ``reps`` launches of one stage
        are captured into a CUDA graph and the replay is timed with CUDA events, so host launch gaps are not in the
        number (an eager loop under-reports the kernel, most of all for the short fp32 launches); best of
        ``rounds`` replays.  The last stage is the variant a step with diagnostics runs (sums fused in)."""
        if self._scratch is None:
            self._scratch = [self.ctx.new_field() for _ in range(3)]
            self._q0 = [x.clone() for x in state]      # distinct from qs, as in stages 2 and 3 of a real step
            self._stage_diag = torch.zeros(2, dtype=torch.float64, device=f"cuda:{self.device}")
        pairs = RK_PAIRS[self.rk]
        ms, sums_us = [0.0] * len(pairs), None
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

        def timed(launch):
            launch()                                                              # warm (lazy module load)
            torch.cuda.synchronize(self.device)
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                for _ in range(reps):
                    launch()
            graph.replay()
            torch.cuda.synchronize(self.device)
            best = float("inf")
            for _ in range(rounds):
                e0.record()
                graph.replay()
                e1.record()
                torch.cuda.synchronize(self.device)
                best = min(best, e0.elapsed_time(e1) / reps)
            del graph
            return best

        for k, (a, b) in enumerate(pairs):
            ms[k] = timed(lambda: self.ctx.rk_stage(self._q0, state, self._scratch, a, b, dt, part=2))
        if self.world == 1:
            # what the fused conservation sums add to a stage launch: of the output (last stage of a step that wants the
            # new state's sums) and of the input (first stage, sums of the entry state)
            a, b = pairs[-1]
            d = self._stage_diag
            out_us = (timed(lambda: self.ctx.rk_stage(self._q0, state, self._scratch, a, b, dt, part=2, diag=d)) - ms[-1]) * 1e3
            a, b = pairs[0]
            in_us = (timed(lambda: self.ctx.rk_stage(self._q0, state, self._scratch, a, b, dt, part=2, diag=d, sums_of_input=True))
                     - ms[0]) * 1e3
            sums_us = {"of_new_state_in_last_stage": out_us, "of_entry_state_in_first_stage": in_us}
        return {"a": [a for a, _ in pairs], "ms": ms, "fused_sums_us": sums_us,
                "kernel": "stage_march_kernel (fused RK stage)", "how": f"{reps} launches per CUDA graph, best of {rounds} replays"}

    def time_halo(self, state, dt, reps=20):
        """Cross-GPU halo numbers (world > 1): ``exchange_us`` = pack + all-to-all + unpack alone, per fill;
        ``step_ms_without_exchange`` = the step with the exchange left out (wrong ghosts, same kernels), from which
        the caller gets the exposed halo time per step."""
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for _ in range(3):
            self._exchange(*state)
        dist.barrier(group=self.group)
        torch.cuda.synchronize(self.device)
        e0.record()
        for _ in range(reps):
            self._exchange(*state)
        e1.record()
        torch.cuda.synchronize(self.device)
        exch_us = e0.elapsed_time(e1) * 1e3 / reps
        # the same step without the exchange: scratch state (the result is not physical), eager launches are replaced
        # by a graph exactly as in step()
        scratch = [x.clone() for x in state]
        self._skip_exchange = True
        saved = (self._graphs, self._eager_calls)
        self._graphs, self._eager_calls = {}, {}
        try:
            for _ in range(4):
                self.step(*scratch, dt)
            dist.barrier(group=self.group)
            torch.cuda.synchronize(self.device)
            e0.record()
            for _ in range(reps):
                self.step(*scratch, dt)
            e1.record()
            torch.cuda.synchronize(self.device)
        finally:
            self._skip_exchange = False
            self._graphs, self._eager_calls = saved
        t = torch.tensor([exch_us, e0.elapsed_time(e1) / reps], device=f"cuda:{self.device}", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)
        return {"exchange_us": float(t[0].item()), "step_ms_without_exchange": float(t[1].item())}

    def time_e2e(self, pinned, dev, dt, nsteps, cells_total):
        """Sharded end-to-end round trip: per step H2D of the shard from pinned memory, halo fill, the step, D2H of
        the new state into pinned memory and of [mass, energy]."""
        back = [torch.empty_like(p).pin_memory() for p in pinned]
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for warm in (True, False):
            if not warm:
                dist.barrier(group=self.group)
                torch.cuda.synchronize()
                e0.record()
            for _ in range(1 if warm else nsteps):
                for d, p in zip(dev, pinned):
                    d.copy_(p, non_blocking=True)
                self.halo_fill(*dev)
                self.step(*dev, dt)
                for b, d in zip(back, dev):
                    b.copy_(d, non_blocking=True)
                self.diagnostics(*dev).cpu()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / nsteps], device=f"cuda:{self.device}", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)
        nbytes = sum(x.numel() * x.element_size() for x in pinned)
        return {"value": cells_total / (float(t.item()) * 1e-3), "unit": "cell-updates/s",
                "h2d_bytes_per_step": nbytes, "d2h_bytes_per_step": nbytes + 16, "ms_per_step": float(t.item()),
                "call": "per rank: H2D of its shard (pinned), halo fill + exchange, step, D2H of the shard and of "
                        "[mass, energy]; bytes are per rank"}
