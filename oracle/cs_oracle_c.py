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
_LIB = _HERE / "_build" / "libcs_oracle.so"
_lib = None
_RK = {"euler": 1, "ssprk2": 2, "ssprk3": 3}
_LIM = {"none": 0, "minmod": 1, "mc": 2}


def lib():
    global _lib
    if _lib is None:
        if not _LIB.exists():
            subprocess.run(["make", "-C", str(_HERE)], check=True)
        h = C.CDLL(str(_LIB))
        dp = C.POINTER(C.c_double)
        h.cso_create.restype = C.c_void_p
        h.cso_create.argtypes = [C.c_int] * 5 + [C.c_double] * 3 + [dp, dp, dp]
        h.cso_destroy.argtypes = [C.c_void_p]
        h.cso_step.argtypes = [C.c_void_p, dp, dp, dp, C.c_double, C.c_int]
        h.cso_diagnostics.argtypes = [C.c_void_p, dp, dp, dp, dp]
        h.cso_halo_fill.argtypes = [C.c_int, C.c_int, dp, dp, dp]
        h.cso_num_threads.restype = C.c_int
        h.cso_set_num_threads.argtypes = [C.c_int]
        _lib = h
    return _lib


def set_threads(n):
    """Number of OpenMP threads for the following calls (the timed baseline uses every core it is allowed)."""
    lib().cso_set_num_threads(int(n))


def _p(a):
    return a.ctypes.data_as(C.POINTER(C.c_double)) if a is not None else None


def _check(a, shape):
    if a.dtype != np.float64 or a.shape != shape or not a.flags.c_contiguous:
        raise ValueError(f"need a C-contiguous float64 array of shape {shape}")


class COracle:
    """Same parameters as :class:`cs_oracle_np.Params`; arrays are face-level ``[6, M, M]`` float64."""

    def __init__(self, N, g=2, a=orc.EARTH_RADIUS, omega=orc.EARTH_OMEGA, grav=orc.EARTH_GRAVITY,
                 axis=(0.0, 0.0, 1.0), limiter="mc", rk="ssprk3", advect_only=False, hs=None):
        self.N, self.g = N, g
        self.shape = (6, N + 2 * g, N + 2 * g)
        axis = np.ascontiguousarray(axis, dtype=np.float64)
        tang = np.ascontiguousarray(orc.centre_tangents(N))
        if hs is not None:
            hs = np.ascontiguousarray(hs, dtype=np.float64)
            _check(hs, self.shape)
        self._h = lib().cso_create(N, g, _LIM[limiter], _RK[rk], int(advect_only), a, omega, grav,
                                   _p(axis), _p(tang), _p(hs))
        self.threads = lib().cso_num_threads()

    def __del__(self):
        if getattr(self, "_h", None):
            lib().cso_destroy(self._h)
            self._h = None

    def step(self, h, u, v, dt, nsteps=1):
        """In place."""
        for a in (h, u, v):
            _check(a, self.shape)
        lib().cso_step(self._h, _p(h), _p(u), _p(v), float(dt), int(nsteps))
        return h, u, v

    def diagnostics(self, h, u, v):
        out = np.zeros(2)
        lib().cso_diagnostics(self._h, _p(h), _p(u), _p(v), _p(out))
        return float(out[0]), float(out[1])


def halo_fill(N, g, h=None, u=None, v=None):
    """In-place ghost fill of a scalar and/or a vector pair."""
    lib().cso_halo_fill(N, g, _p(h), _p(u), _p(v))
