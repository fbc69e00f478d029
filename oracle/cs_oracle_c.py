"""TEST INFRASTRUCTURE -- ctypes loader of the C restatement of the oracle (oracle/cs_oracle.c).

Same header as cs_oracle_np.py applies: parity is unpinned upstream (the reference has no
runnable solver or golden vectors); this C port is pinned to the NumPy oracle by
tests/test_oracle.py.  Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline /
reference arm may import it.
"""
from __future__ import annotations

import ctypes as C
import subprocess
from pathlib import Path

import numpy as np

from . import cs_oracle_np as orc

_HERE = Path(__file__).resolve().parent
_LIBS = {np.dtype(np.float64): _HERE / "_build" / "libcs_oracle.so",
         np.dtype(np.float32): _HERE / "_build" / "libcs_oracle_f32.so"}      # same source, -DCSO_REAL=float
_loaded = {}
_RK = {"euler": 1, "ssprk2": 2, "ssprk3": 3}
_LIM = {"none": 0, "minmod": 1, "mc": 2}


def lib(dtype=np.float64):
    """The C oracle built for ``dtype`` (float64: the parity checker; float32: the fp32 CPU arm of bench.py)."""
    dtype = np.dtype(dtype)
    if dtype not in _loaded:
        path = _LIBS[dtype]
        if not path.exists():
            subprocess.run(["make", "-C", str(_HERE)], check=True)
        h = C.CDLL(str(path))
        assert h.cso_real_bytes() == dtype.itemsize
        vp, dp = C.c_void_p, C.POINTER(C.c_double)
        h.cso_create.restype = C.c_void_p
        h.cso_create.argtypes = [C.c_int] * 5 + [C.c_double] * 3 + [dp, dp, vp]
        h.cso_destroy.argtypes = [C.c_void_p]
        h.cso_step.argtypes = [C.c_void_p, vp, vp, vp, C.c_double, C.c_int]
        h.cso_diagnostics.argtypes = [C.c_void_p, vp, vp, vp, dp]
        h.cso_halo_fill.argtypes = [C.c_int, C.c_int, vp, vp, vp]
        h.cso_num_threads.restype = C.c_int
        h.cso_set_num_threads.argtypes = [C.c_int]
        _loaded[dtype] = h
    return _loaded[dtype]


def set_threads(n, dtype=np.float64):
    """Number of OpenMP threads for the following calls (the timed baseline uses every core it is allowed)."""
    lib(dtype).cso_set_num_threads(int(n))


def _p(a):
    return a.ctypes.data_as(C.POINTER(C.c_double)) if a is not None else None


def _v(a):
    return C.c_void_p(a.ctypes.data) if a is not None else None


def _check(a, shape, dtype=np.float64):
    if a.dtype != np.dtype(dtype) or a.shape != shape or not a.flags.c_contiguous:
        raise ValueError(f"need a C-contiguous {np.dtype(dtype).name} array of shape {shape}")


class COracle:
    """Same parameters as :class:`cs_oracle_np.Params`; arrays are face-level ``[6, M, M]`` of ``dtype``
    (float64: the checker; float32: state, tables and per-cell arithmetic in single precision)."""

    def __init__(self, N, g=2, a=orc.EARTH_RADIUS, omega=orc.EARTH_OMEGA, grav=orc.EARTH_GRAVITY,
                 axis=(0.0, 0.0, 1.0), limiter="mc", rk="ssprk3", advect_only=False, hs=None, dtype=np.float64):
        self.N, self.g, self.dtype = N, g, np.dtype(dtype)
        self._lib = lib(self.dtype)
        self.shape = (6, N + 2 * g, N + 2 * g)
        axis = np.ascontiguousarray(axis, dtype=np.float64)
        tang = np.ascontiguousarray(orc.centre_tangents(N))
        if hs is not None:
            hs = np.ascontiguousarray(hs, dtype=self.dtype)
            _check(hs, self.shape, self.dtype)
        self._h = self._lib.cso_create(N, g, _LIM[limiter], _RK[rk], int(advect_only), a, omega, grav,
                                       _p(axis), _p(tang), _v(hs))
        self.threads = self._lib.cso_num_threads()

    def __del__(self):
        if getattr(self, "_h", None):
            self._lib.cso_destroy(self._h)
            self._h = None

    def step(self, h, u, v, dt, nsteps=1):
        """In place."""
        for a in (h, u, v):
            _check(a, self.shape, self.dtype)
        self._lib.cso_step(self._h, _v(h), _v(u), _v(v), float(dt), int(nsteps))
        return h, u, v

    def diagnostics(self, h, u, v):
        out = np.zeros(2)
        self._lib.cso_diagnostics(self._h, _v(h), _v(u), _v(v), _p(out))
        return float(out[0]), float(out[1])


def halo_fill(N, g, h=None, u=None, v=None):
    """In-place ghost fill of a scalar and/or a vector pair (all of one dtype, float64 or float32)."""
    dtype = next(a.dtype for a in (h, u, v) if a is not None)
    lib(dtype).cso_halo_fill(N, g, _v(h), _v(u), _v(v))
