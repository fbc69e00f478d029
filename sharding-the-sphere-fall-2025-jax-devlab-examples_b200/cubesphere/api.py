"""Host-side mirror of the reference's interface for the hot path.

Same names, argument meaning and error behaviour as the reference snippets
(/root/reference/JAX-DevLab-Examples.py, "PY" below) where they exist:

  ===============================  ===========================================
  reference                        here
  ===============================  ===========================================
  setup_sharding(self)  PY:19-85   :func:`setup_sharding`, :class:`TileMesh`,
                                   :class:`TileSharding`
  create_communication_schedule    re-exported from :mod:`.topology`
    PY:105-139
  apply_operations  PY:143-163     re-exported from :mod:`.topology`
  extract_boundary_data /          :func:`extract_boundary_data`,
  set_ghost_data (called at        :func:`set_ghost_data` (the reference calls
    PY:184-195, never defined)     them but does not ship them)
  exchange_edge_pair  PY:166-197   folded into the one-launch halo kernel; the
                                   schedule is validated, not replayed
  make_halo_exchange  PY:199-246   :func:`make_halo_exchange`
  step(state, dt)  (north_star)    :func:`make_step`
  ===============================  ===========================================

PyTorch is used for device memory and streams only; all arithmetic on the path
runs in the hand-written sm_100a kernels behind the C ABI (no fallback).
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import numpy as np

from . import _binding as B
from . import topology
from .topology import apply_operations, create_communication_schedule, create_tile_schedule  # noqa: F401  (re-export)

try:  # torch is plumbing (device memory, streams); host-only helpers work without it
    import torch
except Exception:  # pragma: no cover
    torch = None


# ----------------------------------------------------------------------------
# device mesh / sharding  (reference PY:19-85)
# ----------------------------------------------------------------------------

@dataclass(frozen=True)
class TileMesh:
    """1-D device mesh with the single axis name 'tiles' (reference PY:77-78)."""
    devices: tuple
    axis_names: tuple = ("tiles",)

    @property
    def size(self):
        return len(self.devices)


@dataclass(frozen=True)
class TileSharding:
    """Block partition of axis 0 (tiles) over the mesh: NamedSharding(mesh, P('tiles')) (PY:79)."""
    mesh: TileMesh
    num_tiles: int
    spec: str = "tiles"

    @property
    def tiles_per_device(self):
        return self.num_tiles // self.mesh.size

    def tile_range(self, rank: int):
        """(tile_begin, tile_count) owned by mesh position ``rank``."""
        if not 0 <= rank < self.mesh.size:
            raise ValueError(f"rank {rank} outside the mesh of {self.mesh.size} device(s)")
        per = self.tiles_per_device
        return rank * per, per

    def owner(self, tile: int):
        return tile // self.tiles_per_device


def setup_sharding(config, verbose=False):
    """Validate the ``parallelization`` block and build the tile mesh / sharding.

    Mirrors reference ``setup_sharding`` (PY:19-85): keys ``device_type``
    (default 'cpu'), ``num_devices`` (6), ``tiles_per_edge`` (1); ``ValueError``
    when there are more devices than tiles (PY:43-48) or the tiles do not divide
    evenly (PY:50-57).  Differences: ``tiles_per_edge > 1`` is implemented here
    (the reference raises NotImplementedError, PY:31-37), and with
    ``device_type='cpu'`` the mesh is a host-side description only -- usable for
    planning and the gloo tests; kernels exist for CUDA devices alone.

    Returns ``(mesh, sharding)``.
    """
    para = config["parallelization"] if "parallelization" in config else config
    device_type = para.get("device_type", "cpu")
    num_devices = int(para.get("num_devices", 6))
    tiles_per_edge = int(para.get("tiles_per_edge", 1))
    if tiles_per_edge < 1:
        raise ValueError(f"tiles_per_edge must be a positive integer, got {tiles_per_edge}")
    if num_devices < 1:
        raise ValueError(f"num_devices must be a positive integer, got {num_devices}")
    n_tiles = topology.num_tiles(tiles_per_edge)
    if num_devices > n_tiles:
        raise ValueError(
            f"{num_devices} devices requested for a sphere of {n_tiles} tiles (tiles_per_edge = {tiles_per_edge}): "
            f"every device needs at least one tile, so at most {n_tiles} devices can be used")
    if n_tiles % num_devices != 0:
        valid = [d for d in range(1, n_tiles + 1) if n_tiles % d == 0]
        raise ValueError(
            f"{n_tiles} tiles (tiles_per_edge = {tiles_per_edge}) cannot be split evenly over {num_devices} devices; "
            f"the tile axis is block-partitioned, so use one of {valid}")
    if device_type == "cpu":
        devices = tuple(("cpu", k) for k in range(num_devices))
    elif device_type == "gpu":
        if torch is None or not torch.cuda.is_available():
            raise RuntimeError("device_type = 'gpu' but no CUDA device is visible")
        avail = torch.cuda.device_count()
        world = 1
        if torch.distributed.is_available() and torch.distributed.is_initialized():
            world = torch.distributed.get_world_size()
        if num_devices > max(avail, world):
            raise ValueError(f"{num_devices} devices requested, {max(avail, world)} CUDA device(s) visible")
        devices = tuple(("cuda", k) for k in range(num_devices))
    else:
        raise ValueError(f"device_type must be 'gpu' or 'cpu', got {device_type!r}")
    mesh = TileMesh(devices)
    sharding = TileSharding(mesh, n_tiles)
    if verbose:
        print(f"  Mesh created: {n_tiles} tiles across {len(devices)} device(s); "
              f"PartitionSpec('tiles') on axis 0, {sharding.tiles_per_device} tile(s) per device")
    return mesh, sharding


# ----------------------------------------------------------------------------
# the reference's two missing helpers (called at PY:184-195)
# ----------------------------------------------------------------------------

def extract_boundary_data(face_field, edge, N, g=1):
    """Interior strip next to ``edge`` of one ``(N+2g, N+2g)`` face, canonical form ``[g, N]``
    (depth from the edge, along-edge index); ``(N,)`` when g == 1 like the reference's data."""
    a = face_field
    if edge == "N":
        s = a[N:g + N, g:g + N].flip(0) if hasattr(a, "flip") else a[N:g + N, g:g + N][::-1]
    elif edge == "S":
        s = a[g:2 * g, g:g + N]
    elif edge == "E":
        t = a[g:g + N, N:g + N].T
        s = t.flip(0) if hasattr(t, "flip") else t[::-1]
    elif edge == "W":
        s = a[g:g + N, g:2 * g].T
    else:
        raise ValueError(f"Unknown edge: {edge}")
    return s[0] if g == 1 else s


def set_ghost_data(face_field, edge, data, N, g=1):
    """Return a copy of the face with the ghost strip on ``edge`` replaced by ``data``
    (canonical ``[g, N]`` or ``(N,)`` when g == 1).  Functional, like ``.at[].set`` (PY:192-195)."""
    out = face_field.clone() if hasattr(face_field, "clone") else face_field.copy()
    G = data.reshape(1, -1) if g == 1 else data
    rev = (lambda x: x.flip(0)) if hasattr(G, "flip") else (lambda x: x[::-1])
    if edge == "N":
        out[g + N:g + N + g, g:g + N] = G
    elif edge == "S":
        out[0:g, g:g + N] = rev(G)
    elif edge == "E":
        out[g:g + N, g + N:g + N + g] = G.T
    elif edge == "W":
        out[g:g + N, 0:g] = rev(G).T
    else:
        raise ValueError(f"Unknown edge: {edge}")
    return out


def validate_schedule(schedule):
    """Check a communication schedule against the cube geometry.

    Every entry must be a real seam with the right op flag, every seam must appear
    exactly once and no face may appear twice in a stage (deck slide 9).  Raises
    ValueError (unknown op: same message as reference PY:163).
    """
    seen = set()
    for stage in schedule:
        faces = []
        for (fa, ea), (fb, eb), ops in stage:
            if ops not in ("N", "T", "R", "TR"):
                raise ValueError(f"seam operation must be one of N, T, R, TR; got {ops!r}")
            nb, ne, _, _ = topology.neighbour(fa, ea)
            if (nb, ne) != (fb, eb):
                raise ValueError(f"({fa},{ea}) <-> ({fb},{eb}) is not a cube seam; ({fa},{ea}) meets ({nb},{ne})")
            if topology.seam_op(fa, ea) != ops:
                raise ValueError(f"seam ({fa},{ea}) <-> ({fb},{eb}) needs op {topology.seam_op(fa, ea)}, got {ops}")
            key = frozenset(((fa, ea), (fb, eb)))
            if key in seen:
                raise ValueError(f"seam ({fa},{ea}) <-> ({fb},{eb}) listed twice")
            seen.add(key)
            faces += [fa, fb]
        if len(set(faces)) != len(faces):
            raise ValueError("a face appears twice in one communication stage")
    if len(seen) != 12:
        raise ValueError(f"schedule covers {len(seen)} of the 12 cube seams")


# ----------------------------------------------------------------------------
# context wrapper
# ----------------------------------------------------------------------------

def _torch_dtype(dtype):
    if torch is None:
        raise RuntimeError("PyTorch is required for device memory and streams")
    if dtype in ("f32", "fp32", "float32", np.float32) or dtype is torch.float32:
        return torch.float32
    if dtype in ("f64", "fp64", "float64", np.float64, float) or dtype is torch.float64:
        return torch.float64
    raise ValueError(f"Unknown dtype: {dtype}")


class Context:
    """Owns one ``cs_ctx`` (C ABI) for a contiguous tile range on one CUDA device."""

    def __init__(self, N, tiles_per_edge=1, halo=2, dtype="f64", rk="ssprk3", limiter="mc", advect_only=False,
                 device=0, tile_begin=0, tile_count=0, kernel="auto", fused_halo=False, radius=0.0, omega=0.0,
                 gravity=0.0, axis=(0.0, 0.0, 1.0), tangents=None, sums_of_entry_state=False, tile_order=None):
        self._lib = B.lib()
        self.dtype = _torch_dtype(dtype)
        if rk not in B.RK_IDS:
            raise ValueError(f"Unknown rk scheme: {rk}")
        if limiter not in B.LIMITER_IDS:
            raise ValueError(f"Unknown limiter: {limiter}")
        if kernel not in B.KERNEL_IDS:
            raise ValueError(f"Unknown kernel variant: {kernel}")
        cfg = B.CsConfig()
        cfg.abi_version = B.CS_ABI_VERSION
        cfg.N, cfg.tiles_per_edge, cfg.halo = int(N), int(tiles_per_edge), int(halo)
        cfg.dtype = B.CS_F64 if self.dtype is torch.float64 else B.CS_F32
        cfg.rk, cfg.limiter = B.RK_IDS[rk], B.LIMITER_IDS[limiter]
        cfg.advect_only = int(bool(advect_only))
        cfg.device, cfg.tile_begin, cfg.tile_count = int(device), int(tile_begin), int(tile_count)
        cfg.kernel, cfg.fused_halo = B.KERNEL_IDS[kernel], int(bool(fused_halo))
        cfg.sums_of_entry_state = int(bool(sums_of_entry_state))
        cfg.radius, cfg.omega, cfg.gravity = float(radius), float(omega), float(gravity)
        for k in range(3):
            cfg.axis[k] = float(axis[k])
        self._tangents = None
        if tangents is not None:
            self._tangents = np.ascontiguousarray(tangents, dtype=np.float64)
            if self._tangents.shape != (int(N),):
                raise ValueError(f"tangents must have shape ({N},)")
            cfg.tangents = self._tangents.ctypes.data_as(C.POINTER(C.c_double))
        # tile_order: position along axis 0 -> tile id (a permutation of range(6 p^2)); tile_begin / tile_count then
        # count positions.  None = identity (the reference's plain block order).
        n_all = topology.num_tiles(int(tiles_per_edge))
        self.tile_order = np.arange(n_all, dtype=np.int32)
        if tile_order is not None:
            self.tile_order = np.ascontiguousarray(tile_order, dtype=np.int32)
            if self.tile_order.shape != (n_all,):
                raise ValueError(f"tile_order must have {n_all} entries")
            cfg.tile_order = self.tile_order.ctypes.data_as(C.POINTER(C.c_int32))
        handle = C.c_void_p()
        B.check(self._lib.cs_ctx_create(C.byref(cfg), C.byref(handle)))
        self._h = handle
        self.N, self.p, self.g = int(N), int(tiles_per_edge), int(halo)
        self.n = self.N // self.p
        self.m = self.n + 2 * self.g
        self.device = int(device)
        self.tile_begin = int(tile_begin)
        self.tile_count = int(tile_count) or topology.num_tiles(self.p)
        self.tiles = self.tile_order[self.tile_begin:self.tile_begin + self.tile_count].copy()   # tile ids owned, in order
        self.shape = (self.tile_count, self.m, self.m)

    # -- plumbing -------------------------------------------------------------
    def close(self):
        if getattr(self, "_h", None):
            self._lib.cs_ctx_destroy(self._h)
            self._h = None

    def __del__(self):  # pragma: no cover
        try:
            self.close()
        except Exception:
            pass

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def _ptr(self, t, name="field"):
        if not isinstance(t, torch.Tensor) or not t.is_cuda:
            raise ValueError(f"{name} must be a CUDA tensor (there is no CPU path)")
        if t.device.index != self.device:
            raise ValueError(f"{name} lives on cuda:{t.device.index}, the context on cuda:{self.device}")
        if t.dtype is not self.dtype:
            raise ValueError(f"{name} has dtype {t.dtype}, the context computes in {self.dtype}")
        if tuple(t.shape) != self.shape:
            raise ValueError(f"{name} has shape {tuple(t.shape)}, expected {self.shape}")
        if not t.is_contiguous():
            raise ValueError(f"{name} must be contiguous")
        return C.c_void_p(t.data_ptr())

    def _check(self, rc):
        B.check(rc, self._h)

    def new_field(self):
        return torch.zeros(self.shape, dtype=self.dtype, device=f"cuda:{self.device}")

    @property
    def launches(self):
        return int(self._lib.cs_launch_count(self._h))

    def set_option(self, name, value):
        """Tuning knob of this context by name (see ``cs_ctx_set_option``)."""
        self._check(self._lib.cs_ctx_set_option(self._h, name.encode(), int(value)))

    def stage_kernel_info(self):
        """(variant name of the last stage launch, number of launches an 'auto' context had to send to the
        one-thread-per-cell kernel because the layout was not vectorisable)."""
        last, falls = C.c_int32(0), C.c_int64(0)
        self._check(self._lib.cs_stage_kernel_info(self._h, C.byref(last), C.byref(falls)))
        names = {v: k for k, v in B.KERNEL_IDS.items()}
        return names.get(last.value, str(last.value)), int(falls.value)

    # -- halo -----------------------------------------------------------------
    def halo_fill_scalar(self, field):
        self._check(self._lib.cs_halo_fill_scalar(self._h, self._ptr(field), self._stream()))
        return field

    def halo_fill_vector(self, u, v):
        self._check(self._lib.cs_halo_fill_vector(self._h, self._ptr(u, "u"), self._ptr(v, "v"), self._stream()))
        return u, v

    def halo_fill_state(self, h, u, v):
        self._check(self._lib.cs_halo_fill_state(self._h, self._ptr(h, "h"), self._ptr(u, "u"), self._ptr(v, "v"),
                                                 self._stream()))
        return h, u, v

    def halo_table(self):
        """(dst_offset, src_offset, code) of every ghost cell; local cells first (the order a
        rotation table must follow), remote ones (src_offset == -1) after."""
        n = C.c_size_t()
        self._check(self._lib.cs_halo_table_info(self._h, 0, None, None, None, C.byref(n)))
        dst = np.empty(n.value, dtype=np.int64)
        src = np.empty(n.value, dtype=np.int64)
        code = np.empty(n.value, dtype=np.int32)
        self._check(self._lib.cs_halo_table_info(
            self._h, n.value, dst.ctypes.data_as(C.POINTER(C.c_int64)), src.ctypes.data_as(C.POINTER(C.c_int64)),
            code.ctypes.data_as(C.POINTER(C.c_int32)), C.byref(n)))
        return dst, src, code

    def set_rotation(self, rot):
        """Install (or clear with None) a position-dependent 2x2 rotation ``[n_local, 2, 2]``."""
        if rot is None:
            self._check(self._lib.cs_halo_set_rotation(self._h, None, 0))
            return
        rot = np.ascontiguousarray(rot, dtype=np.float64).reshape(-1, 4)
        self._check(self._lib.cs_halo_set_rotation(self._h, rot.ctypes.data_as(C.POINTER(C.c_double)), rot.shape[0]))

    def set_halo_interpolation(self, tables):
        """Install (or clear with None) second-order ghosts: ``tables`` = ``geometry.seam_interp_tables(N, g, p)``
        (keys dst, src [n, 4], w [n, 4], rot [n, 2, 2], tangents).  Fill ghosts again afterwards; a topography set
        earlier has to be set again too (its ghosts and the metric change).  See ``cs_halo_set_interp``."""
        if tables is None:
            self._check(self._lib.cs_halo_set_interp(self._h, 0, None, None, None, None, None))
            return
        dst = np.ascontiguousarray(tables["dst"], dtype=np.int64)
        src = np.ascontiguousarray(tables["src"], dtype=np.int64).reshape(-1, 4)
        w = np.ascontiguousarray(tables["w"], dtype=np.float64).reshape(-1, 4)
        rot = np.ascontiguousarray(tables["rot"], dtype=np.float64).reshape(-1, 4)
        tang = tables.get("tangents")
        tp = None
        if tang is not None:
            tang = np.ascontiguousarray(tang, dtype=np.float64)
            if tang.shape != (self.N + 2 * self.g,):
                raise ValueError(f"tangents must have shape ({self.N + 2 * self.g},)")
            tp = tang.ctypes.data_as(C.POINTER(C.c_double))
        if not (dst.shape[0] == src.shape[0] == w.shape[0] == rot.shape[0]):
            raise ValueError("dst, src, w, rot must have one row per ghost cell")
        i64, f64 = C.POINTER(C.c_int64), C.POINTER(C.c_double)
        self._check(self._lib.cs_halo_set_interp(self._h, dst.shape[0], dst.ctypes.data_as(i64), src.ctypes.data_as(i64),
                                                 w.ctypes.data_as(f64), rot.ctypes.data_as(f64), tp))

    def halo_interp_sharded(self, plan):
        """Second-order ghosts on a shard: ``plan`` = ``partition.InterpExchangePlan`` of this rank.  Replaces
        ``halo_plan``; afterwards ``halo_push`` + ``halo_pull`` produce every ghost cell.  See ``cs_halo_interp_sharded``."""
        i64, f64 = C.POINTER(C.c_int64), C.POINTER(C.c_double)
        n = len(plan.send_counts)
        sc = (C.c_int64 * n)(*plan.send_counts)
        rc = (C.c_int64 * n)(*plan.recv_counts)
        ss = np.ascontiguousarray(np.concatenate(plan.send_src) if n else np.empty(0), dtype=np.int64)
        dst = np.ascontiguousarray(plan.dst, dtype=np.int64)
        src = np.ascontiguousarray(plan.src, dtype=np.int64).reshape(-1, 4)
        w = np.ascontiguousarray(plan.w, dtype=np.float64).reshape(-1, 4)
        rot = np.ascontiguousarray(plan.rot, dtype=np.float64).reshape(-1, 4)
        tang = np.ascontiguousarray(plan.tangents, dtype=np.float64)
        self._check(self._lib.cs_halo_interp_sharded(self._h, n, sc, ss.ctypes.data_as(i64) if ss.size else None, rc, dst.shape[0],
                                                     dst.ctypes.data_as(i64), src.ctypes.data_as(i64), w.ctypes.data_as(f64),
                                                     rot.ctypes.data_as(f64), tang.ctypes.data_as(f64)))

    # -- step -----------------------------------------------------------------
    def set_topography(self, hs):
        if hs is None:
            self._check(self._lib.cs_set_topography(self._h, None, self._stream()))
        else:
            self._check(self._lib.cs_set_topography(self._h, self._ptr(hs, "hs"), self._stream()))

    def _triple(self, q, name):
        return (C.c_void_p * 3)(*[self._ptr(t, f"{name}[{k}]") for k, t in enumerate(q)])

    def rk_stage(self, q0, qs, qout, a, b, dt, diag=None, part=None, sums_of_input=False):
        q0p = self._triple(q0, "q0") if q0 is not None else (C.c_void_p * 3)(None, None, None)
        if part is None:
            dp = C.c_void_p(diag.data_ptr()) if diag is not None else None
            self._check(self._lib.cs_rk_stage(self._h, q0p, self._triple(qs, "qs"), self._triple(qout, "qout"),
                                              float(a), float(b), float(dt), dp, self._stream()))
        else:
            # a split stage: ``diag`` receives the sums over the cells of this part only, of the new values or
            # (sums_of_input) of qs at those cells
            dp = C.c_void_p(diag.data_ptr()) if diag is not None else None
            which = B.CS_SUMS_OF_INPUT if sums_of_input else B.CS_SUMS_OF_OUTPUT
            self._check(self._lib.cs_rk_stage_part_sums(self._h, q0p, self._triple(qs, "qs"), self._triple(qout, "qout"),
                                                        float(a), float(b), float(dt), int(part), which, dp, self._stream()))
        return qout

    def step(self, h, u, v, dt, diag=None):
        dp = C.c_void_p(diag.data_ptr()) if diag is not None else None
        self._check(self._lib.cs_step(self._h, self._ptr(h, "h"), self._ptr(u, "u"), self._ptr(v, "v"), float(dt), dp,
                                      self._stream()))
        return h, u, v

    def diagnostics(self, h, u, v, out=None):
        """Device tensor ``[mass, energy]`` (fp64)."""
        if out is None:
            out = torch.empty(2, dtype=torch.float64, device=f"cuda:{self.device}")
        self._check(self._lib.cs_diagnostics(self._h, self._ptr(h, "h"), self._ptr(u, "u"), self._ptr(v, "v"),
                                             C.c_void_p(out.data_ptr()), self._stream()))
        return out

    def scalar_diffuse(self, T, Tout, kappa_dt, source=None, dt_source=0.0):
        """One explicit diffusion step of a scalar tracer (the reference's heat-diffusion demo): ``Tout`` gets
        ``T + kappa_dt L(T) + dt_source source`` and filled ghosts.  ``T`` must be ghost-filled."""
        sp = self._ptr(source, "source") if source is not None else None
        self._check(self._lib.cs_scalar_diffuse(self._h, self._ptr(T, "T"), self._ptr(Tout, "Tout"), sp, float(kappa_dt),
                                                float(dt_source), self._stream()))
        return Tout

    def _host_ptrs(self, h, u, v):
        ptrs = []
        np_dtype = np.float64 if self.dtype is torch.float64 else np.float32
        for name, a in (("h", h), ("u", u), ("v", v)):
            if torch is not None and isinstance(a, torch.Tensor):
                if a.is_cuda or a.dtype is not self.dtype or tuple(a.shape) != self.shape or not a.is_contiguous():
                    raise ValueError(f"{name}: need a contiguous CPU tensor of shape {self.shape} and dtype {self.dtype}")
                ptrs.append(C.c_void_p(a.data_ptr()))
            else:
                if not isinstance(a, np.ndarray) or a.dtype != np_dtype or a.shape != self.shape or not a.flags.c_contiguous:
                    raise ValueError(f"{name}: need a C-contiguous array of shape {self.shape} and dtype {np_dtype.__name__}")
                ptrs.append(C.c_void_p(a.ctypes.data))
        return ptrs

    def step_host(self, h, u, v, dt, nsteps=1, copy_back=True, diag=True):
        """Host-buffer entry: NumPy arrays (or pinned CPU tensors) in the padded layout, updated in place."""
        ptrs = self._host_ptrs(h, u, v)
        out = (C.c_double * 2)()
        self._check(self._lib.cs_step_host(self._h, ptrs[0], ptrs[1], ptrs[2], float(dt), int(nsteps),
                                           out if diag else None, int(bool(copy_back))))
        return (out[0], out[1]) if diag else None

    def step_host_batch(self, members, dt, nsteps=1, copy_back=True, diag=True):
        """``members``: sequence of (h, u, v) host states (an ensemble).  Uploads, steps and downloads overlap
        across members (two device slots); page-locked arrays are needed for that.  Returns [members, 2]
        mass / energy (or None)."""
        M = len(members)
        cols = list(zip(*(self._host_ptrs(*m) for m in members))) if M else [(), (), ()]
        arrs = [(C.c_void_p * M)(*col) for col in cols]
        out = (C.c_double * (2 * max(M, 1)))()
        self._check(self._lib.cs_step_host_batch(self._h, M, arrs[0], arrs[1], arrs[2], float(dt), int(nsteps),
                                                 out if diag else None, int(bool(copy_back))))
        return np.array(out[:2 * M], dtype=np.float64).reshape(M, 2) if diag else None

    # -- sharded exchange -----------------------------------------------------
    def halo_plan(self, tile_begins, tile_counts):
        nb = len(tile_begins)
        tb = (C.c_int32 * nb)(*tile_begins)
        tcnt = (C.c_int32 * nb)(*tile_counts)
        send = (C.c_int64 * nb)()
        recv = (C.c_int64 * nb)()
        self._check(self._lib.cs_halo_plan(self._h, nb, tb, tcnt, send, recv))
        return list(send), list(recv)

    def halo_pack(self, h, u, v, sendbuf, nfields=3):
        up = self._ptr(u, "u") if u is not None else None
        vp = self._ptr(v, "v") if v is not None else None
        self._check(self._lib.cs_halo_pack(self._h, self._ptr(h, "h"), up, vp, nfields,
                                           C.c_void_p(sendbuf.data_ptr()), self._stream()))

    def halo_unpack(self, recvbuf, h, u, v, nfields=3):
        up = self._ptr(u, "u") if u is not None else None
        vp = self._ptr(v, "v") if v is not None else None
        self._check(self._lib.cs_halo_unpack(self._h, C.c_void_p(recvbuf.data_ptr()), self._ptr(h, "h"), up, vp,
                                             nfields, self._stream()))

    # one-sided exchange over CUDA IPC (GPUs of one box)
    def halo_peer_alloc(self):
        """This context's IPC handle block (bytes) for its receive buffers and flags; gather all ranks' blocks."""
        buf = C.create_string_buffer(B.CS_PEER_HANDLE_BYTES)
        self._check(self._lib.cs_halo_peer_alloc(self._h, buf, B.CS_PEER_HANDLE_BYTES))
        return buf.raw

    def halo_peer_connect(self, my_rank, all_handles, remote_offsets):
        blob = b"".join(all_handles)
        n = len(all_handles)
        if len(blob) != n * B.CS_PEER_HANDLE_BYTES or len(remote_offsets) != n:
            raise ValueError("need one handle block and one offset per rank")
        off = (C.c_int64 * n)(*[int(x) for x in remote_offsets])
        self._check(self._lib.cs_halo_peer_connect(self._h, n, int(my_rank), C.c_char_p(blob), off))

    def halo_peer_pointers(self):
        """(recv0, recv1, flags) device addresses of this context's exchange buffers: for contexts of one process."""
        out = (C.c_void_p * 3)()
        self._check(self._lib.cs_halo_peer_pointers(self._h, out))
        return [int(x or 0) for x in out]

    def halo_peer_connect_pointers(self, my_rank, all_pointers, remote_offsets):
        n = len(all_pointers)
        flat = (C.c_void_p * (3 * n))(*[p for trip in all_pointers for p in trip])
        off = (C.c_int64 * n)(*[int(x) for x in remote_offsets])
        self._check(self._lib.cs_halo_peer_connect_pointers(self._h, n, int(my_rank), flat, off))

    def halo_peer_timeouts(self):
        """Number of pulls that gave up waiting for a peer (0 in a healthy run)."""
        n = C.c_uint(0)
        self._check(self._lib.cs_halo_peer_timeouts(self._h, C.byref(n)))
        return int(n.value)

    def halo_push(self, h, u, v, nfields=3):
        up = self._ptr(u, "u") if u is not None else None
        vp = self._ptr(v, "v") if v is not None else None
        self._check(self._lib.cs_halo_push(self._h, self._ptr(h, "h"), up, vp, nfields, self._stream()))

    def halo_pull(self, h, u, v, nfields=3):
        up = self._ptr(u, "u") if u is not None else None
        vp = self._ptr(v, "v") if v is not None else None
        self._check(self._lib.cs_halo_pull(self._h, self._ptr(h, "h"), up, vp, nfields, self._stream()))


# ----------------------------------------------------------------------------
# the reference's factories
# ----------------------------------------------------------------------------

def make_halo_exchange(schedule, N, *, halo=1, tiles_per_edge=1, dtype="f64", device=0, inplace=False):
    """Factory for the compiled halo exchange, like reference ``make_halo_exchange(schedule, N)``
    (PY:199-246): all static arguments (faces, edges, ops, N) are frozen at creation and the
    returned callable maps ``field_ghosts (6 p^2, N/p + 2g, N/p + 2g) -> field_ghosts``.

    The 12 swaps of the schedule are validated against the cube geometry and then executed as ONE
    kernel launch (each ghost cell has a single writer, so the 4-stage ordering is not needed).
    The callable is functional by default (returns a new tensor, like ``.at[].set``); pass
    ``inplace=True`` to fill the given tensor.  ``exchange.vector(u, v)`` fills a vector field,
    ``exchange.state(h, u, v)`` all three in one launch.
    """
    validate_schedule(schedule)
    ctx = Context(N, tiles_per_edge=tiles_per_edge, halo=halo, dtype=dtype, device=device)

    def cubesphere_halo_exchange(field_ghosts):
        out = field_ghosts if inplace else field_ghosts.clone()
        return ctx.halo_fill_scalar(out)

    def vector(u, v):
        if not inplace:
            u, v = u.clone(), v.clone()
        return ctx.halo_fill_vector(u, v)

    def state(h, u, v):
        if not inplace:
            h, u, v = h.clone(), u.clone(), v.clone()
        return ctx.halo_fill_state(h, u, v)

    cubesphere_halo_exchange.vector = vector
    cubesphere_halo_exchange.state = state
    cubesphere_halo_exchange.context = ctx
    return cubesphere_halo_exchange


def make_step(N, *, tiles_per_edge=1, halo=2, dtype="f64", rk="ssprk3", limiter="mc", advect_only=False,
              device=0, hs=None, axis=(0.0, 0.0, 1.0), kernel="auto", fused_halo=False, tangents=None, **phys):
    """Return ``step(state, dt) -> state`` for a single-device run (all 6 p^2 tiles on ``device``).

    ``state`` is a dict ``{'h', 'u', 'v'}`` of halo-filled CUDA tensors ``[6 p^2, n + 2g, n + 2g]``;
    it is advanced in place and returned (the pytree surface named by BASELINE.json's north_star).
    ``step.diagnostics(state)`` returns the device tensor ``[mass, energy]``.
    """
    ctx = Context(N, tiles_per_edge=tiles_per_edge, halo=halo, dtype=dtype, rk=rk, limiter=limiter,
                  advect_only=advect_only, device=device, axis=axis, kernel=kernel, fused_halo=fused_halo,
                  tangents=tangents, **phys)
    if hs is not None:
        ctx.set_topography(hs)

    def step(state, dt):
        ctx.step(state["h"], state["u"], state["v"], dt)
        return state

    step.context = ctx
    step.diagnostics = lambda state: ctx.diagnostics(state["h"], state["u"], state["v"])
    return step
