"""cs_diagnostics_noalloc: the context-free, allocation-free mass / energy reduction.

It exists for callers that cannot create a cs_ctx because they cannot allocate device memory
(a plugin handler inside a framework that has reserved the GPU for its own allocator, a monitoring
hook on a latency-critical thread).  Its only scratch is a fixed static array inside the library,
shared by every call on the device, so the checks below run it twice in a row on different data:
the second call must not be disturbed by what the first one left in the array.  Results are
compared with the context-based cs_diagnostics AND, independently of any product kernel, with the
NumPy oracle's diagnostics of the same numbers, within the tolerances used elsewhere in this suite
(mass 1e-13, energy 1e-12, relative).  The file name sorts last on purpose.
"""
import ctypes as C

import numpy as np
import pytest

torch = pytest.importorskip("torch")

pytestmark = pytest.mark.gpu

from cubesphere import _binding as B  # noqa: E402
from cubesphere import api, cases  # noqa: E402
from oracle import cs_oracle_np as orc  # noqa: E402


def _noalloc(lib, dtype, N, p, g, h, u, v, hs=None, tile_begin=0, tile_count=0):
    out = torch.full((2,), float("nan"), dtype=torch.float64, device="cuda")
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = lib.cs_diagnostics_noalloc(
        B.CS_F64 if dtype is torch.float64 else B.CS_F32, N, p, g, tile_begin, tile_count, 0.0, 0.0,
        C.c_void_p(h.data_ptr()), C.c_void_p(u.data_ptr()), C.c_void_p(v.data_ptr()),
        C.c_void_p(hs.data_ptr()) if hs is not None else None, C.c_void_p(out.data_ptr()), 0, stream)
    return rc, out


def _state(name, N, p, g, dtype):
    case = cases.make_case(name, N)
    par = orc.Params(N, g, axis=case.axis, hs=case.hs)
    faces = orc.halo_fill_state(orc.pad(case.h, g), orc.pad(case.u1, g), orc.pad(case.u2, g), g)
    dev = [torch.from_numpy(orc.faces_to_tiles(x, p, g)).to(dtype).cuda() for x in faces]
    hs = torch.from_numpy(orc.faces_to_tiles(par.hs, p, g)).to(dtype).cuda() if case.hs is not None else None
    return case, dev, hs


def _oracle_diag(case, N, g, dtype):
    """[mass, energy] by the NumPy oracle of exactly the numbers the device holds (rounded to fp32 when it is)."""
    nd = np.float64 if dtype is torch.float64 else np.float32
    rnd = lambda x: x.astype(nd).astype(np.float64)
    par = orc.Params(N, g, axis=case.axis, hs=None if case.hs is None else rnd(case.hs))
    faces = [rnd(x) for x in orc.halo_fill_state(orc.pad(case.h, g), orc.pad(case.u1, g), orc.pad(case.u2, g), g)]
    return np.array(orc.diagnostics(par, *faces))


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
def test_two_calls_in_a_row_match_context_diagnostics(built_library, dtype):
    lib = built_library
    g = 2
    # call 1: large grid (many block partials in the shared scratch); call 2: small grid with topography
    for name, N, p in (("tc2", 192, 1), ("tc5", 48, 2)):
        case, dev, hs = _state(name, N, p, g, dtype)
        ctx = api.Context(N, tiles_per_edge=p, halo=g, dtype=dtype, axis=case.axis)
        if hs is not None:
            ctx.set_topography(hs)
        want = ctx.diagnostics(*dev).cpu().numpy()
        rc, out = _noalloc(lib, dtype, N, p, g, *dev, hs=hs)
        assert rc == B.CS_OK, lib.cs_last_error(None)
        got = out.cpu().numpy()
        assert abs(got[0] - want[0]) <= 1e-13 * abs(want[0]), (name, got, want)
        assert abs(got[1] - want[1]) <= 1e-12 * abs(want[1]), (name, got, want)
        ref = _oracle_diag(case, N, g, dtype)             # the independent CPU statement, no product code involved
        assert abs(got[0] - ref[0]) <= 1e-13 * abs(ref[0]), (name, got, ref)
        assert abs(got[1] - ref[1]) <= 1e-12 * abs(ref[1]), (name, got, ref)
        # same call again: bit-identical (fixed summation order, stale scratch ignored)
        rc, again = _noalloc(lib, dtype, N, p, g, *dev, hs=hs)
        assert rc == B.CS_OK and torch.equal(again, out)
        ctx.close()


def test_shard_range_and_limits(built_library):
    lib = built_library
    g, N, p = 2, 48, 2
    case, dev, _ = _state("tc2", N, p, g, torch.float64)
    ctx = api.Context(N, tiles_per_edge=p, halo=g, axis=case.axis)
    whole = ctx.diagnostics(*dev).cpu().numpy()
    ctx.close()
    parts = np.zeros(2)
    for b in range(0, 24, 6):                            # four shards of six tiles sum to the whole sphere
        shard = [x[b:b + 6].contiguous() for x in dev]
        rc, out = _noalloc(lib, torch.float64, N, p, g, *shard, tile_begin=b, tile_count=6)
        assert rc == B.CS_OK
        parts += out.cpu().numpy()
    assert np.all(np.abs(parts - whole) <= 1e-13 * np.abs(whole))
    # beyond the fixed scratch: a clear error, nothing launched (pointers are never dereferenced)
    rc, _ = _noalloc(lib, torch.float64, 3072, 1, g, *dev)
    assert rc == B.CS_E_INVALID and b"exceed the fixed scratch" in lib.cs_last_error(None)
    rc, _ = _noalloc(lib, torch.float64, 50, 4, g, *dev)
    assert rc == B.CS_E_INVALID


def test_callers_current_device_is_left_alone(built_library):
    """The callers this entry point is for sit inside another framework's handler thread: the device that thread
    had selected must still be selected when the call returns."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two visible GPUs to tell a restored device from an untouched one")
    lib = built_library
    g, N = 2, 24
    case, dev, _ = _state("tc2", N, 1, g, torch.float64)          # state on cuda:0
    torch.cuda.synchronize(0)
    torch.cuda.set_device(1)
    try:
        out = torch.zeros(2, dtype=torch.float64, device="cuda:0")
        rc = lib.cs_diagnostics_noalloc(B.CS_F64, N, 1, g, 0, 0, 0.0, 0.0, C.c_void_p(dev[0].data_ptr()),
                                        C.c_void_p(dev[1].data_ptr()), C.c_void_p(dev[2].data_ptr()), None,
                                        C.c_void_p(out.data_ptr()), 0, None)
        assert rc == B.CS_OK
        assert torch.cuda.current_device() == 1
    finally:
        torch.cuda.set_device(0)
    torch.cuda.synchronize(0)
    ref = _oracle_diag(case, N, g, torch.float64)
    assert np.all(np.abs(out.cpu().numpy() - ref) <= 1e-12 * np.abs(ref))
