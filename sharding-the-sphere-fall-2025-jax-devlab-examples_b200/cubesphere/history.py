"""History files in the zarr v2 directory layout, written without the zarr package (host side).

The reference names zarr history files (deck slides 4 and 6: "history -> zarr") but ships no writer, and zarr is not
installed in this image.  The v2 directory store is simple enough to write directly: a group is a directory with a
``.zgroup`` file, an array is a sub-directory with a ``.zarray`` JSON (shape, chunks, dtype, order, compressor) and one
file per chunk named by the chunk's indices joined with dots; attributes live in ``.zattrs``.  The stores written here
are uncompressed (``"compressor": null``), C-ordered, little-endian, so ``zarr.open(path)`` / ``xarray.open_zarr(path)``
on any machine that has those packages reads them as they are, and this module reads them back with NumPy alone.

Layout of a history::

    <path>/.zgroup, .zattrs                      format, N, tiles_per_edge, halo, case, dtype
    <path>/h, u, v      [record, tile, y, x]     chunks [1, 1, m, m]: ONE FILE PER (record, tile)
    <path>/time, mass, energy   [record] f8      chunks [256]
    <path>/step                 [record] i8      chunks [256]
    <path>/lon, lat     [tile, y, x] f8          degrees at the cell centres (ghost cells: the tile's own chart extended),
                                                 written once at creation; the fields name them as CF "coordinates"

One chunk per (record, tile) is what makes the store fit the sharded solver: a rank writes the chunk files of the tiles
it owns, straight from its own shard -- no gather -- and one rank (``write_meta=True``) appends the scalars and, last of
all, publishes the new record count by replacing the ``.zarray`` files (written to a temporary name first, so a reader
or a crashed run never sees a record whose chunks are incomplete).
"""
from __future__ import annotations

import json
import os
from pathlib import Path

import numpy as np

FORMAT = "cubesphere-b200-history-v1"
FIELDS = ("h", "u", "v")
SCALARS = (("time", "<f8"), ("step", "<i8"), ("mass", "<f8"), ("energy", "<f8"))
SCALAR_CHUNK = 256


def _to_numpy(a):
    if hasattr(a, "detach"):            # torch tensor (CPU or CUDA)
        a = a.detach().cpu().numpy()
    return np.ascontiguousarray(a)


def _write_json(path: Path, obj):
    tmp = path.with_name(path.name + ".tmp")
    tmp.write_text(json.dumps(obj, indent=1, sort_keys=True))
    os.replace(tmp, path)               # atomic on POSIX: a reader sees the old or the new file, never half of one


def tile_lonlat(N, p, g):
    """Longitude / latitude in degrees of every cell of the padded tile layout: two arrays [6 p^2, m, m]."""
    from . import geometry as geo
    n = N // p
    frames = np.asarray(geo.FACE_FRAMES, dtype=np.float64)
    lon = np.empty((6 * p * p, n + 2 * g, n + 2 * g))
    lat = np.empty_like(lon)
    ext = np.tan(-0.25 * np.pi + (np.arange(-g, N + g) + 0.5) * (0.5 * np.pi / N))     # the chart continued past the edges
    ext[g:g + N] = geo.centre_tangents(N)                                              # interior: the solver's own table
    for t in range(6 * p * p):
        f, r = divmod(t, p * p)
        ty, tx = divmod(r, p)
        X, Y = np.meshgrid(ext[tx * n:tx * n + n + 2 * g], ext[ty * n:ty * n + n + 2 * g], indexing="xy")
        P = frames[f, 0] + X[..., None] * frames[f, 1] + Y[..., None] * frames[f, 2]
        P /= np.linalg.norm(P, axis=-1, keepdims=True)
        lo, la = geo.lonlat(P)
        lon[t], lat[t] = np.degrees(lo), np.degrees(la)
    return lon, lat


def _zarray(shape, chunks, dtype):
    return {"zarr_format": 2, "shape": [int(x) for x in shape], "chunks": [int(x) for x in chunks], "dtype": dtype,
            "compressor": None, "fill_value": 0 if dtype[1] == "i" else 0.0, "filters": None, "order": "C"}


class ZarrHistory:
    """Append-only history of the solver state.  ``create`` makes an empty store, ``open`` an existing one."""

    def __init__(self, path, attrs):
        self.path = Path(path)
        self.attrs = attrs
        self.N, self.p, self.g = attrs["N"], attrs["tiles_per_edge"], attrs["halo"]
        self.m = self.N // self.p + 2 * self.g
        self.tiles = 6 * self.p * self.p
        self.dtype = np.dtype(attrs["dtype"])

    # ---- creation / opening -----------------------------------------------------------------------------------
    @classmethod
    def create(cls, path, *, N, tiles_per_edge=1, halo=2, dtype="float64", case=None, extra=None, coordinates=True):
        path = Path(path)
        if (path / ".zgroup").exists():
            raise FileExistsError(f"{path} already holds a zarr group")
        if N < 1 or tiles_per_edge < 1 or N % tiles_per_edge:
            raise ValueError(f"N = {N} must be a positive multiple of tiles_per_edge = {tiles_per_edge}")
        dt = np.dtype(dtype)
        if dt not in (np.dtype(np.float32), np.dtype(np.float64)):
            raise ValueError(f"dtype must be float32 or float64, got {dtype}")
        attrs = dict(format=FORMAT, N=int(N), tiles_per_edge=int(tiles_per_edge), halo=int(halo), case=case,
                     dtype=dt.name, layout="[record, tile, y + ghosts, x + ghosts]; tile = face * p^2 + ty * p + tx",
                     extra=extra or {})
        self = cls(path, attrs)
        path.mkdir(parents=True, exist_ok=True)
        _write_json(path / ".zgroup", {"zarr_format": 2})
        _write_json(path / ".zattrs", attrs)
        for name in FIELDS:
            (path / name).mkdir(exist_ok=True)
            _write_json(path / name / ".zattrs", dict({"_ARRAY_DIMENSIONS": ["record", "tile", "y", "x"]},
                                                      **({"coordinates": "lon lat"} if coordinates else {})))
            _write_json(path / name / ".zarray", _zarray((0, self.tiles, self.m, self.m), (1, 1, self.m, self.m),
                                                          dt.newbyteorder("<").str))
        for name, code in SCALARS:
            (path / name).mkdir(exist_ok=True)
            _write_json(path / name / ".zattrs", {"_ARRAY_DIMENSIONS": ["record"]})
            _write_json(path / name / ".zarray", _zarray((0,), (SCALAR_CHUNK,), code))
        if coordinates:
            for name, a, long_name in zip(("lon", "lat"), tile_lonlat(N, tiles_per_edge, halo), ("longitude", "latitude")):
                (path / name).mkdir(exist_ok=True)
                _write_json(path / name / ".zattrs", {"_ARRAY_DIMENSIONS": ["tile", "y", "x"], "units": "degrees_" +
                                                      ("east" if name == "lon" else "north"), "long_name": long_name})
                for t in range(self.tiles):
                    a[t].astype("<f8").tofile(path / name / f"{t}.0.0")
                _write_json(path / name / ".zarray", _zarray(a.shape, (1, self.m, self.m), "<f8"))
        return self

    @classmethod
    def open(cls, path):
        path = Path(path)
        try:
            attrs = json.loads((path / ".zattrs").read_text())
        except FileNotFoundError:
            raise FileNotFoundError(f"{path}: not a history store (no .zattrs)") from None
        if attrs.get("format") != FORMAT:
            raise ValueError(f"{path}: not a {FORMAT} store")
        return cls(path, attrs)

    def __len__(self):
        return int(json.loads((self.path / "h" / ".zarray").read_text())["shape"][0])

    # ---- writing ---------------------------------------------------------------------------------------------
    def append(self, state, *, step, time, mass=float("nan"), energy=float("nan"), tile_ids=None, write_meta=True,
               record=None):
        """Write one record.  ``state`` = {'h', 'u', 'v'} in the padded tile layout ``[tiles_here, m, m]`` (NumPy arrays or
        torch tensors).  Whole sphere, one writer: the defaults.  Sharded: every rank passes the global ids of the tiles
        it holds (``tile_ids``, in the order of its axis 0) and the record index it writes (``record``; all ranks agree,
        e.g. ``len(history)`` read before the first of them publishes); exactly one passes ``write_meta=True`` and it must
        do so AFTER the others have finished (a barrier in the caller).  Returns the index of the record."""
        k = len(self) if record is None else int(record)
        arrays = {name: _to_numpy(state[name]) for name in FIELDS}
        ids = list(range(self.tiles)) if tile_ids is None else [int(t) for t in tile_ids]
        for name, a in arrays.items():
            if a.shape != (len(ids), self.m, self.m):
                raise ValueError(f"{name}: expected shape {(len(ids), self.m, self.m)}, got {a.shape}")
            if a.dtype != self.dtype:
                raise ValueError(f"{name}: the store holds {self.dtype.name}, got {a.dtype}")
        if any(t < 0 or t >= self.tiles for t in ids) or len(set(ids)) != len(ids):
            raise ValueError("tile_ids must be distinct tile numbers of this sphere")
        le = self.dtype.newbyteorder("<")
        for name, a in arrays.items():
            for row, t in enumerate(ids):
                a[row].astype(le, copy=False).tofile(self.path / name / f"{k}.{t}.0.0")
        if write_meta:
            for (name, code), value in zip(SCALARS, (time, step, mass, energy)):
                c, r = divmod(k, SCALAR_CHUNK)
                f = self.path / name / str(c)
                chunk = np.fromfile(f, dtype=code) if f.exists() else np.zeros(SCALAR_CHUNK, dtype=code)
                chunk[r] = value
                tmp = f.with_name(f.name + ".tmp")
                chunk.tofile(tmp)
                os.replace(tmp, f)
            # publish: the scalars first, the fields last -- __len__ reads h's shape
            for name, _ in SCALARS:
                meta = json.loads((self.path / name / ".zarray").read_text())
                meta["shape"] = [k + 1]
                _write_json(self.path / name / ".zarray", meta)
            for name in reversed(FIELDS):
                meta = json.loads((self.path / name / ".zarray").read_text())
                meta["shape"][0] = k + 1
                _write_json(self.path / name / ".zarray", meta)
        return k

    # ---- reading ---------------------------------------------------------------------------------------------
    def scalars(self):
        """{'time', 'step', 'mass', 'energy'}: 1-D arrays over the records."""
        n, out = len(self), {}
        for name, code in SCALARS:
            parts = [np.fromfile(self.path / name / str(c), dtype=code) for c in range((n + SCALAR_CHUNK - 1) // SCALAR_CHUNK)]
            out[name] = (np.concatenate(parts) if parts else np.zeros(0, dtype=code))[:n]
        return out

    def read(self, k=-1, tile_ids=None):
        """Record ``k`` (negative: from the end) as ``(state, meta)``; ``tile_ids`` selects tiles (a rank's own shard)."""
        n = len(self)
        if k < 0:
            k += n
        if not 0 <= k < n:
            raise IndexError(f"record {k} of {n}")
        ids = list(range(self.tiles)) if tile_ids is None else [int(t) for t in tile_ids]
        le = self.dtype.newbyteorder("<")
        state = {}
        for name in FIELDS:
            out = np.empty((len(ids), self.m, self.m), dtype=self.dtype)
            for row, t in enumerate(ids):
                f = self.path / name / f"{k}.{t}.0.0"
                # a chunk that was never written reads as the fill value (zarr semantics)
                out[row] = np.fromfile(f, dtype=le).reshape(self.m, self.m) if f.exists() else 0
            state[name] = out
        sc = self.scalars()
        meta = dict(self.attrs, record=k, step=int(sc["step"][k]), time=float(sc["time"][k]), mass=float(sc["mass"][k]),
                    energy=float(sc["energy"][k]))
        return state, meta


def write_sharded(hist, state, *, tile_ids, rank, barrier, step, time, mass=float("nan"), energy=float("nan")):
    """One record written by all ranks of a job (every rank has opened the same store on a shared file system and calls
    this with its own shard): each rank writes the chunk files of the tiles it owns, rank 0 adds the scalars and
    publishes the record once everybody is done.  ``barrier`` is the job's barrier (``torch.distributed.barrier``, or a
    no-op for one rank).  Returns the record index."""
    record = len(hist)                      # the same on every rank: nothing of this record is published yet ...
    barrier()                               # ... and nobody publishes before everybody has read it
    if rank != 0:
        hist.append(state, step=step, time=time, tile_ids=tile_ids, write_meta=False, record=record)
    barrier()
    if rank == 0:
        hist.append(state, step=step, time=time, mass=mass, energy=energy, tile_ids=tile_ids, record=record)
    barrier()
    return record
