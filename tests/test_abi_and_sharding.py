"""CPU-side checks of the boundary: the shared library loads and exports every symbol the header
declares (no compute calls without a GPU), fails loudly without CUDA, and the sharded exchange plan
works over gloo with world_size 2."""
import ctypes
import os
import re
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest

from cubesphere import _binding, partition
from oracle import cs_oracle_np as orc

ROOT = Path(__file__).resolve().parent.parent


def test_library_exports_every_declared_symbol(built_library):
    header = (ROOT / "include" / "cubesphere.h").read_text()
    declared = sorted(set(re.findall(r"^CS_API\s+[\w\s\*]+?\b(cs_\w+)\(", header, flags=re.M)))
    assert declared == _binding.exported_symbols()            # binding and header list the same entry points
    for name in declared:
        assert hasattr(built_library, name), name
    assert built_library.cs_abi_version() == _binding.CS_ABI_VERSION


def test_config_struct_matches_header():
    header = (ROOT / "include" / "cubesphere.h").read_text()
    body = re.search(r"typedef struct cs_config \{(.*?)\} cs_config;", header, flags=re.S).group(1)
    fields = re.findall(r"^\s*(?:const\s+)?(?:int32_t|double|double\*)\s*\*?\s*(\w+)", body, flags=re.M)
    assert fields == [f[0] for f in _binding.CsConfig._fields_]


def test_header_is_plain_c(tmp_path):
    """The boundary is a C ABI: the header must compile as C99 with nothing but the C library."""
    import shutil
    import subprocess
    gcc = "/usr/bin/gcc" if Path("/usr/bin/gcc").exists() else shutil.which("gcc")
    if not gcc:
        pytest.skip("no C compiler")
    src = tmp_path / "t.c"
    src.write_text('#include "cubesphere.h"\nint main(void) { struct cs_config c; c.N = 1; return cs_abi_version() == c.N; }\n')
    inc = Path(__file__).resolve().parents[1] / "include"
    r = subprocess.run([gcc, "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-fsyntax-only", f"-I{inc}", str(src)],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr


def test_no_cpu_fallback(built_library):
    """Without a CUDA device context creation fails with CS_E_CUDA and a clear message."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    cfg = _binding.CsConfig()
    cfg.abi_version, cfg.N, cfg.tiles_per_edge, cfg.halo, cfg.dtype = 1, 24, 1, 2, _binding.CS_F64
    cfg.rk, cfg.limiter = 3, 2
    h = ctypes.c_void_p()
    rc = built_library.cs_ctx_create(ctypes.byref(cfg), ctypes.byref(h))
    assert rc == _binding.CS_E_CUDA and not h.value
    assert b"no CPU fallback" in built_library.cs_last_error(None)
    with pytest.raises(RuntimeError):
        _binding.check(rc)
    cfg.N = 25
    cfg.tiles_per_edge = 2                                     # argument errors come first, as ValueError
    rc = built_library.cs_ctx_create(ctypes.byref(cfg), ctypes.byref(h))
    assert rc == _binding.CS_E_INVALID
    with pytest.raises(ValueError):
        _binding.check(rc)


def test_missing_library_is_loud(monkeypatch, tmp_path):
    monkeypatch.setenv("CUBESPHERE_B200_LIB", str(tmp_path / "nope.so"))
    monkeypatch.setattr(_binding, "_lib", None)
    with pytest.raises(RuntimeError, match="no CPU or PyTorch fallback"):
        _binding.lib()
    monkeypatch.setattr(_binding, "_lib", None)


def test_product_never_imports_oracle():
    pkg = ROOT / "sharding-the-sphere-fall-2025-jax-devlab-examples_b200"
    for f in list(pkg.rglob("*.py")) + list(pkg.rglob("*.cu")) + list(pkg.rglob("*.cuh")):
        text = f.read_text()
        assert "oracle" not in text.replace("the oracle", "").replace("NumPy oracle", "").lower() or \
            not re.search(r"^\s*(from|import)\s+oracle|#include\s+\".*oracle", text, flags=re.M), f


@pytest.mark.parametrize("tile_map", ["block", "locality", "shuffled"])
@pytest.mark.parametrize("N,p,world", [(8, 1, 2), (8, 2, 4), (12, 2, 8), (8, 1, 6)])
def test_exchange_plan_serial_emulation(N, p, world, tile_map):
    """All ranks emulated in one process: local fill + pack/exchange/unpack == face-level oracle fill, whatever the
    order of the tiles along the sharded axis (the reference's block order, the compact blocks of locality_order, or
    any permutation)."""
    order = {"block": np.arange(6 * p * p), "locality": partition.locality_order(p, world),
             "shuffled": np.random.default_rng(3).permutation(6 * p * p)}[tile_map]
    g = 2
    rng = np.random.default_rng(7)
    M = N + 2 * g
    faces = [rng.standard_normal((6, M, M)) for _ in range(3)]
    ref = [orc.faces_to_tiles(x, p, g) for x in orc.halo_fill_state(*faces, g)]
    n = N // p
    tiles = []
    for x in faces:
        t = orc.faces_to_tiles(x, p, g)
        w = np.full_like(t, np.nan)
        w[:, g:g + n, g:g + n] = t[:, g:g + n, g:g + n]
        tiles.append(w)
    ranges = partition.block_ranges(6 * p * p, world)
    plans = [partition.ExchangePlan(N, p, g, r, ranges, order) for r in range(world)]
    shards = [[np.ascontiguousarray(x[order[b:b + c]]) for x in tiles] for b, c in ranges]
    for r in range(world):
        for q in range(world):
            assert plans[r].send_counts[q] == plans[q].recv_counts[r]
    packed = [pl.pack(sh) for pl, sh in zip(plans, shards)]
    for r, (pl, sh) in enumerate(zip(plans, shards)):
        pl.fill_local(sh)
        pl.unpack([packed[q][r] for q in range(world)], sh)
        b, c = ranges[r]
        for got, want in zip(sh, ref):
            assert np.array_equal(got, want[order[b:b + c]])


def test_locality_order_cuts_fewer_edges():
    """partition.locality_order keeps the reference's contract (equal contiguous blocks of axis 0, PY:77-79) and only
    reorders the tiles: 24 tiles on 8 devices -> the three tiles around each cube vertex (24 of 48 tile edges cross
    devices, 36 in block order; at most 3 neighbour devices), on 4 devices 16 instead of 26."""
    ident = np.arange(24)
    assert partition.cut_edges(ident, 2, 8) == 36 and partition.cut_edges(ident, 2, 4) == 26
    o8, o4 = partition.locality_order(2, 8), partition.locality_order(2, 4)
    for o in (o8, o4):
        assert sorted(o.tolist()) == list(range(24))
    assert partition.cut_edges(o8, 2, 8) == 24 and partition.cut_edges(o4, 2, 4) == 16
    assert np.array_equal(o8, partition.locality_order(2, 8))              # deterministic: every rank computes the same
    owner = np.empty(24, dtype=int)
    owner[o8] = np.arange(24) // 3
    nb = [set() for _ in range(8)]
    for a, b in partition.tile_adjacency(2):
        if owner[a] != owner[b]:
            nb[owner[a]].add(owner[b]), nb[owner[b]].add(owner[a])
    assert max(len(x) for x in nb) == 3
    assert len(partition.tile_adjacency(2)) == 48 and len(partition.tile_adjacency(1)) == 12
    assert np.array_equal(partition.locality_order(1, 1), np.arange(6))
    with pytest.raises(ValueError):
        partition.locality_order(1, 4)


_GLOO_WORKER = r"""
import os, sys
sys.path.insert(0, sys.argv[1]); sys.path.insert(0, sys.argv[2])
import numpy as np, torch, torch.distributed as dist
from cubesphere import partition
from oracle import cs_oracle_np as orc
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
N, p, g = 12, 2, 2
rng = np.random.default_rng(11)
M = N + 2 * g
faces = [rng.standard_normal((6, M, M)) for _ in range(3)]
ref = [orc.faces_to_tiles(x, p, g) for x in orc.halo_fill_state(*faces, g)]
n = N // p
ranges = partition.block_ranges(6 * p * p, world)
b, c = ranges[rank]
shard = []
for x in faces:
    t = orc.faces_to_tiles(x, p, g)[b:b + c]
    w = np.full_like(t, np.nan); w[:, g:g + n, g:g + n] = t[:, g:g + n, g:g + n]
    shard.append(np.ascontiguousarray(w))
plan = partition.ExchangePlan(N, p, g, rank, ranges)
send = [torch.from_numpy(np.ascontiguousarray(x)) for x in plan.pack(shard)]
recv = [torch.empty(3, cnt, dtype=torch.float64) for cnt in plan.recv_counts]
reqs = []
for q in range(world):
    if q == rank: continue
    reqs.append(dist.isend(send[q], q)); reqs.append(dist.irecv(recv[q], q))
for r in reqs: r.wait()
plan.fill_local(shard)
plan.unpack([x.numpy() for x in recv], shard)
ok = all(np.array_equal(a, w[b:b + c]) for a, w in zip(shard, ref))
# the history store, written by both ranks: each its own tiles' chunks, rank 0 publishes (two records)
from pathlib import Path
from cubesphere import history
store = Path(sys.argv[3]) / "run.zarr"
if rank == 0:
    history.ZarrHistory.create(store, N=N, tiles_per_edge=p, halo=g)
dist.barrier()
hist = history.ZarrHistory.open(store)
state = dict(zip("huv", shard))
for k in range(2):
    rec = history.write_sharded(hist, {n: x + k for n, x in state.items()}, tile_ids=range(b, b + c), rank=rank,
                                barrier=dist.barrier, step=5 * k, time=0.5 * k, mass=1.0, energy=2.0)
    ok = ok and rec == k
whole, meta = hist.read(-1)
ok = ok and len(hist) == 2 and meta["step"] == 5 and all(np.array_equal(whole[n], w + 1) for n, w in zip("huv", ref))
t = torch.tensor([int(ok)]); dist.all_reduce(t, op=dist.ReduceOp.MIN)
dist.destroy_process_group()
sys.exit(0 if int(t.item()) == 1 else 1)
"""


def test_exchange_over_gloo_world2(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(_GLOO_WORKER)
    pkg = ROOT / "sharding-the-sphere-fall-2025-jax-devlab-examples_b200"
    env = dict(os.environ, OMP_NUM_THREADS="1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", "29731", str(script), str(pkg), str(ROOT), str(tmp_path)]
    res = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-2000:]


def test_bench_reference_arm_prints_one_contract_line():
    """`bench.py --impl reference` (the C/OpenMP oracle on the host cores) needs no GPU: exactly one JSON line on
    stdout, carrying the contract's keys."""
    import json
    import subprocess
    import sys
    from pathlib import Path
    root = Path(__file__).resolve().parents[1]
    r = subprocess.run([sys.executable, str(root / "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0"],
                       capture_output=True, text=True, timeout=600, cwd=root)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1, r.stdout
    d = json.loads(lines[0])
    for key in ("impl", "metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
                "vs_baseline", "dtype", "data", "config", "cpu_baseline", "e2e"):
        assert key in d, key
    assert d["impl"] == "reference" and d["value"] > 0 and d["unit"] == "cell-updates/s" and d["steps"] == 1
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] >= 1
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0


def _build_abi_demo(tmp_path):
    import shutil
    gcc = "/usr/bin/gcc" if Path("/usr/bin/gcc").exists() else shutil.which("gcc")
    if not gcc:
        pytest.skip("no C compiler")
    libdir = ROOT / "sharding-the-sphere-fall-2025-jax-devlab-examples_b200" / "lib"
    exe = tmp_path / "abi_demo"
    r = subprocess.run([gcc, "-std=c99", "-Wall", "-Wextra", "-pedantic", f"-I{ROOT / 'include'}", str(ROOT / "examples" / "abi_demo.c"),
                        f"-L{libdir}", "-lcubesphere_b200", f"-Wl,-rpath,{libdir}", "-lm", "-o", str(exe)],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return exe


def test_c_caller_links_and_fails_loudly_without_a_gpu(built_library, tmp_path):
    """examples/abi_demo.c is a complete plain-C caller of the header: it must compile and link against the shared
    library with a C compiler alone, and without a CUDA device report the library's 'no CPU fallback' error."""
    import torch
    exe = _build_abi_demo(tmp_path)
    if torch.cuda.is_available():
        pytest.skip("a GPU is present: covered by the gpu-marked test")
    r = subprocess.run([str(exe), "24", "2"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 2 and "no CPU fallback" in r.stderr, (r.returncode, r.stderr)


@pytest.mark.gpu
def test_c_caller_runs_on_the_gpu(built_library, tmp_path):
    """The same program on a GPU: a resting layer stays exactly at rest and its mass is 4 pi a^2 h to O(1/N^2)."""
    exe = _build_abi_demo(tmp_path)
    r = subprocess.run([str(exe), "48", "10"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, (r.stdout, r.stderr)
    assert "max |h - h0| = 0.000e+00" in r.stdout and "max |u| = 0.000e+00" in r.stdout, r.stdout


@pytest.mark.parametrize("tile_map", ["block", "locality"])
@pytest.mark.parametrize("N,p,world", [(12, 1, 2), (12, 2, 4), (16, 2, 8), (12, 1, 6)])
def test_interp_exchange_plan_serial_emulation(N, p, world, tile_map):
    """Second-order ghosts on a sharded sphere: every rank packs the raw cells its peers' rows read, receives its own and
    interpolates -- the edge ghosts equal those of the oracle study's whole-sphere construction to round-off, and every
    ghost cell equals the product's whole-sphere tables applied on one device."""
    from cubesphere import geometry as geo
    from oracle import seam_study
    g = 2
    order = np.arange(6 * p * p) if tile_map == "block" else partition.locality_order(p, world)
    rng = np.random.default_rng(9)
    M = N + 2 * g
    faces = [rng.standard_normal((6, M, M)) for _ in range(3)]
    tiles = [orc.faces_to_tiles(x, p, g) for x in faces]
    n = N // p
    m = n + 2 * g
    # reference 1: the whole-sphere product tables in NumPy
    tab = geo.seam_interp_tables(N, g, p)
    want = [x.copy() for x in tiles]
    flat = [x.reshape(-1) for x in tiles]
    gh, ga, gb = (np.sum(tab["w"] * f[tab["src"]], -1) for f in flat)
    want[0].reshape(-1)[tab["dst"]] = gh
    want[1].reshape(-1)[tab["dst"]] = tab["rot"][:, 0, 0] * ga + tab["rot"][:, 0, 1] * gb
    want[2].reshape(-1)[tab["dst"]] = tab["rot"][:, 1, 0] * ga + tab["rot"][:, 1, 1] * gb
    ranges = partition.block_ranges(6 * p * p, world)
    plans = [partition.InterpExchangePlan(N, p, g, r, ranges, order) for r in range(world)]
    shards = [[np.ascontiguousarray(x[order[b:b + c]]) for x in tiles] for b, c in ranges]
    for r in range(world):
        for q in range(world):
            assert plans[r].send_counts[q] == plans[q].recv_counts[r]
            assert plans[r].remote_offsets[q] == sum(plans[q].recv_counts[:r])
    packed = [pl.pack(sh) for pl, sh in zip(plans, shards)]
    for r, (pl, sh) in enumerate(zip(plans, shards)):
        pl.apply([packed[q][r] for q in range(world)], sh)
        b, c = ranges[r]
        for got, w in zip(sh, want):
            assert np.abs(got - w[order[b:b + c]]).max() <= 1e-13
    # reference 2 (p = 1 only: the study works on whole faces): the oracle's independent construction, edge strips
    if p == 1:
        halo = seam_study.SeamHalo(N, g, "interp")
        ref = halo(*(x.copy() for x in faces))
        for r, sh in enumerate(shards):
            b, c = ranges[r]
            for got, w in zip(sh, ref):
                wt = orc.faces_to_tiles(w, p, g)[order[b:b + c]].reshape(-1)
                d = np.isin(np.arange(wt.size), plans[r].dst) & np.isin((np.arange(wt.size) + b * m * m) if tile_map == "block" else np.full(wt.size, -1), halo.dst)
                if tile_map == "block":
                    assert np.abs(got.reshape(-1)[d] - wt[d]).max() <= 1e-13


def test_integration_notes_name_every_entry_point():
    """INTEGRATION.md is the maintainer's map of the ABI: an entry point added to the header without a word there fails."""
    import re
    root = Path(__file__).resolve().parent.parent
    header = (root / "include" / "cubesphere.h").read_text()
    notes = (root / "INTEGRATION.md").read_text()
    names = set(re.findall(r"CS_API\s+[\w\s\*]+?\b(cs_\w+)\s*\(", header))
    assert len(names) > 30
    missing = sorted(n for n in names if n not in notes)
    assert not missing, missing
