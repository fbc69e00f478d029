"""History files: the zarr v2 directory layout written by cubesphere.history, checked against the
format's rules with an independent reader written from the specification (zarr itself is not installed here)."""
import itertools
import json

import numpy as np
import pytest

from cubesphere import history


def zarr_v2_read(path):
    """Minimal zarr v2 reader: whole array from .zarray + chunk files (uncompressed, C order, '.' separator)."""
    meta = json.loads((path / ".zarray").read_text())
    assert meta["zarr_format"] == 2 and meta["compressor"] is None and meta["filters"] is None and meta["order"] == "C"
    shape, chunks, dt = meta["shape"], meta["chunks"], np.dtype(meta["dtype"])
    out = np.full(shape, meta["fill_value"], dtype=dt)
    grid = [range((s + c - 1) // c) for s, c in zip(shape, chunks)]
    for idx in itertools.product(*grid):
        f = path / ".".join(str(i) for i in idx)
        if not f.exists():
            continue                                    # missing chunk = fill value
        raw = np.fromfile(f, dtype=dt)
        assert raw.size == int(np.prod(chunks)), "a chunk file always holds a whole chunk, edge chunks included"
        block = raw.reshape(chunks)
        sel = tuple(slice(i * c, min((i + 1) * c, s)) for i, c, s in zip(idx, chunks, shape))
        out[sel] = block[tuple(slice(0, s.stop - s.start) for s in sel)]
    return out


def make_state(rng, tiles, m, dtype):
    return {k: rng.standard_normal((tiles, m, m)).astype(dtype) for k in ("h", "u", "v")}


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_round_trip_and_zarr_layout(tmp_path, dtype):
    rng = np.random.default_rng(1)
    N, p, g = 8, 2, 2
    hist = history.ZarrHistory.create(tmp_path / "run.zarr", N=N, tiles_per_edge=p, halo=g, dtype=dtype, case="tc5")
    assert len(hist) == 0
    recs = [make_state(rng, 24, 8, dtype) for _ in range(3)]
    for k, st in enumerate(recs):
        assert hist.append(st, step=10 * k, time=100.0 * k, mass=1.0 + k, energy=2.0 + k) == k
    assert len(hist) == 3
    # through the module
    again = history.ZarrHistory.open(tmp_path / "run.zarr")
    st, meta = again.read(-1)
    assert all(np.array_equal(st[k], recs[2][k]) for k in "huv") and st["h"].dtype == dtype
    assert meta["step"] == 20 and meta["time"] == 200.0 and meta["mass"] == 3.0 and meta["case"] == "tc5" and meta["N"] == N
    own, _ = again.read(1, tile_ids=[5, 3])
    assert np.array_equal(own["u"], recs[1]["u"][[5, 3]])
    # through the format: an independent reader sees the same arrays
    root = tmp_path / "run.zarr"
    assert json.loads((root / ".zgroup").read_text()) == {"zarr_format": 2}
    for name in "huv":
        arr = zarr_v2_read(root / name)
        assert arr.shape == (3, 24, 8, 8) and arr.dtype == np.dtype(dtype)
        assert np.array_equal(arr, np.stack([r[name] for r in recs]))
        assert json.loads((root / name / ".zattrs").read_text())["_ARRAY_DIMENSIONS"] == ["record", "tile", "y", "x"]
    assert np.array_equal(zarr_v2_read(root / "step"), [0, 10, 20])
    assert np.array_equal(zarr_v2_read(root / "energy"), [2.0, 3.0, 4.0])
    with pytest.raises(IndexError):
        again.read(3)


def test_sharded_writers_publish_last(tmp_path):
    """Two ranks write the chunks of their own tiles; the record exists only once the metadata writer has published it."""
    rng = np.random.default_rng(2)
    hist = history.ZarrHistory.create(tmp_path / "h", N=4, tiles_per_edge=1, halo=2, dtype="float64")
    full = make_state(rng, 6, 8, np.float64)
    order = [4, 0, 2, 5, 1, 3]                                     # a locality order: rank 0 holds tiles 4, 0, 2
    k = len(hist)
    r1 = history.ZarrHistory.open(tmp_path / "h")
    r1.append({n: full[n][order[3:]] for n in "huv"}, step=7, time=1.5, tile_ids=order[3:], write_meta=False, record=k)
    assert len(hist) == 0                                          # not published yet
    hist.append({n: full[n][order[:3]] for n in "huv"}, step=7, time=1.5, mass=3.0, tile_ids=order[:3], record=k)
    assert len(hist) == 1
    st, meta = hist.read(0)
    assert all(np.array_equal(st[n], full[n]) for n in "huv") and meta["step"] == 7
    assert not list((tmp_path / "h").rglob("*.tmp"))


def test_scalar_chunks_and_errors(tmp_path, monkeypatch):
    monkeypatch.setattr(history, "SCALAR_CHUNK", 4)
    rng = np.random.default_rng(3)
    hist = history.ZarrHistory.create(tmp_path / "h", N=2, halo=1, dtype="float32")
    st = make_state(rng, 6, 4, np.float32)
    for k in range(6):
        hist.append(st, step=k, time=0.5 * k)
    sc = hist.scalars()
    assert np.array_equal(sc["step"], np.arange(6)) and np.array_equal(sc["time"], 0.5 * np.arange(6))
    assert np.array_equal(zarr_v2_read(tmp_path / "h" / "time"), 0.5 * np.arange(6))      # two chunk files of 4
    assert np.isnan(sc["mass"]).all()
    with pytest.raises(ValueError):
        hist.append(make_state(rng, 6, 4, np.float64), step=0, time=0.0)                 # wrong dtype
    with pytest.raises(ValueError):
        hist.append(make_state(rng, 5, 4, np.float32), step=0, time=0.0)                 # wrong shape
    with pytest.raises(ValueError):
        hist.append({n: st[n][:2] for n in "huv"}, step=0, time=0.0, tile_ids=[1, 1])    # a tile twice
    with pytest.raises(FileExistsError):
        history.ZarrHistory.create(tmp_path / "h", N=2, halo=1)
    with pytest.raises(FileNotFoundError):
        history.ZarrHistory.open(tmp_path / "nothing")
    with pytest.raises(ValueError):
        history.ZarrHistory.create(tmp_path / "x", N=5, tiles_per_edge=2)


def test_coordinates_are_written_once(tmp_path):
    """lon / lat of every stored cell (CF 'coordinates' of the fields): interior cells are where the initial conditions
    were sampled, and a ghost cell's coordinates are those of the neighbour cell it mirrors only to first order (the
    tile's own chart extended), which is what a plot of the padded arrays wants."""
    from cubesphere import geometry as geo
    N, p, g = 8, 2, 2
    history.ZarrHistory.create(tmp_path / "h", N=N, tiles_per_edge=p, halo=g)
    root = tmp_path / "h"
    lon, lat = zarr_v2_read(root / "lon"), zarr_v2_read(root / "lat")
    assert lon.shape == lat.shape == (24, 8, 8)
    assert json.loads((root / "h" / ".zattrs").read_text())["coordinates"] == "lon lat"
    patch = geo.Patch(N, p)
    lo, la = geo.lonlat(patch.points())
    assert np.allclose(lon[:, g:-g, g:-g], np.degrees(lo), atol=1e-12)
    assert np.allclose(lat[:, g:-g, g:-g], np.degrees(la), atol=1e-12)
    assert np.all((lon >= 0) & (lon < 360)) and np.all(np.abs(lat) <= 90)
    # ghosts continue the tile's own chart: across an inner tile edge they ARE the neighbour tile's interior cells
    assert np.allclose(lat[0][g:-g, -g:], lat[1][g:-g, g:2 * g], atol=1e-12)
    # ... and across a cube edge they lie beyond it: one cell further along the same coordinate line, not wrapped around
    east = lon[1][g:-g, -g - 1:]                                    # last interior column and the two ghosts of tile 1
    step = np.diff(np.unwrap(np.radians(east), axis=1), axis=1)
    assert np.all(step[:, 0] * step[:, 1] > 0) and np.all(np.abs(step) < np.radians(360.0 / (4 * N) * 2))
    # face 4 / 5 hold the poles (whichever the frames put there): some tile reaches |lat| > 80
    assert np.abs(lat).max() > 80
    history.ZarrHistory.create(tmp_path / "bare", N=N, coordinates=False)
    assert not (tmp_path / "bare" / "lon").exists()
