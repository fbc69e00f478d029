"""ctypes binding of libcubesphere_b200.so (the C ABI declared in include/cubesphere.h).

There is no fallback of any kind: if the shared library is missing or fails to
load, importing :func:`lib` raises, and every compute call needs a CUDA device.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

PKG_ROOT = Path(__file__).resolve().parent.parent
LIB_PATH = PKG_ROOT / "lib" / "libcubesphere_b200.so"

CS_ABI_VERSION = 1
CS_OK, CS_E_INVALID, CS_E_UNSUPPORTED, CS_E_CUDA, CS_E_NOMEM = 0, -1, -2, -3, -4
CS_F32, CS_F64 = 0, 1
CS_SUMS_OF_OUTPUT, CS_SUMS_OF_INPUT = 0, 1
CS_PEER_HANDLE_BYTES = 192
RK_IDS = {"euler": 1, "ssprk2": 2, "ssprk3": 3}
LIMITER_IDS = {"none": 0, "minmod": 1, "mc": 2}
KERNEL_IDS = {"auto": 0, "naive": 1, "march": 2, "tma": 3}


class CsConfig(C.Structure):
    """Mirror of ``struct cs_config`` (include/cubesphere.h)."""
    _fields_ = [
        ("abi_version", C.c_int32), ("N", C.c_int32), ("tiles_per_edge", C.c_int32), ("halo", C.c_int32),
        ("dtype", C.c_int32), ("rk", C.c_int32), ("limiter", C.c_int32), ("advect_only", C.c_int32),
        ("device", C.c_int32), ("tile_begin", C.c_int32), ("tile_count", C.c_int32), ("kernel", C.c_int32),
        ("fused_halo", C.c_int32), ("sums_of_entry_state", C.c_int32), ("tile_order", C.POINTER(C.c_int32)),
        ("radius", C.c_double), ("omega", C.c_double), ("gravity", C.c_double), ("axis", C.c_double * 3),
        ("tangents", C.POINTER(C.c_double)),
    ]


_VP = C.c_void_p
_PROTOTYPES = {
    "cs_abi_version": (C.c_int, []),
    "cs_last_error": (C.c_char_p, [_VP]),
    "cs_ctx_create": (C.c_int, [C.POINTER(CsConfig), C.POINTER(_VP)]),
    "cs_ctx_destroy": (C.c_int, [_VP]),
    "cs_halo_fill_scalar": (C.c_int, [_VP, _VP, _VP]),
    "cs_halo_fill_vector": (C.c_int, [_VP, _VP, _VP, _VP]),
    "cs_halo_fill_state": (C.c_int, [_VP, _VP, _VP, _VP, _VP]),
    "cs_halo_set_rotation": (C.c_int, [_VP, C.POINTER(C.c_double), C.c_size_t]),
    "cs_halo_set_interp": (C.c_int, [_VP, C.c_size_t, C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_double),
                                     C.POINTER(C.c_double), C.POINTER(C.c_double)]),
    "cs_halo_interp_sharded": (C.c_int, [_VP, C.c_int, C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.c_size_t,
                                         C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_double), C.POINTER(C.c_double),
                                         C.POINTER(C.c_double)]),
    "cs_halo_table_info": (C.c_int, [_VP, C.c_size_t, C.POINTER(C.c_int64), C.POINTER(C.c_int64),
                                     C.POINTER(C.c_int32), C.POINTER(C.c_size_t)]),
    "cs_set_topography": (C.c_int, [_VP, _VP, _VP]),
    "cs_rk_stage": (C.c_int, [_VP, _VP * 3, _VP * 3, _VP * 3, C.c_double, C.c_double, C.c_double, _VP, _VP]),
    "cs_rk_stage_part": (C.c_int, [_VP, _VP * 3, _VP * 3, _VP * 3, C.c_double, C.c_double, C.c_double, C.c_int, _VP]),
    "cs_rk_stage_part_sums": (C.c_int, [_VP, _VP * 3, _VP * 3, _VP * 3, C.c_double, C.c_double, C.c_double, C.c_int, C.c_int, _VP, _VP]),
    "cs_step": (C.c_int, [_VP, _VP, _VP, _VP, C.c_double, _VP, _VP]),
    "cs_diagnostics": (C.c_int, [_VP, _VP, _VP, _VP, _VP, _VP]),
    "cs_step_host": (C.c_int, [_VP, _VP, _VP, _VP, C.c_double, C.c_int, C.POINTER(C.c_double), C.c_int]),
    "cs_step_host_batch": (C.c_int, [_VP, C.c_int, C.POINTER(_VP), C.POINTER(_VP), C.POINTER(_VP), C.c_double, C.c_int,
                                     C.POINTER(C.c_double), C.c_int]),
    "cs_halo_plan": (C.c_int, [_VP, C.c_int, C.POINTER(C.c_int32), C.POINTER(C.c_int32),
                               C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    "cs_halo_pack": (C.c_int, [_VP, _VP, _VP, _VP, C.c_int, _VP, _VP]),
    "cs_halo_unpack": (C.c_int, [_VP, _VP, _VP, _VP, _VP, C.c_int, _VP]),
    "cs_halo_peer_alloc": (C.c_int, [_VP, _VP, C.c_size_t]),
    "cs_halo_peer_connect": (C.c_int, [_VP, C.c_int, C.c_int, _VP, C.POINTER(C.c_int64)]),
    "cs_halo_peer_pointers": (C.c_int, [_VP, C.POINTER(_VP)]),
    "cs_halo_peer_connect_pointers": (C.c_int, [_VP, C.c_int, C.c_int, C.POINTER(_VP), C.POINTER(C.c_int64)]),
    "cs_halo_peer_timeouts": (C.c_int, [_VP, C.POINTER(C.c_uint)]),
    "cs_halo_push": (C.c_int, [_VP, _VP, _VP, _VP, C.c_int, _VP]),
    "cs_halo_pull": (C.c_int, [_VP, _VP, _VP, _VP, C.c_int, _VP]),
    "cs_scalar_diffuse": (C.c_int, [_VP, _VP, _VP, _VP, C.c_double, C.c_double, _VP]),
    "cs_ctx_set_option": (C.c_int, [_VP, C.c_char_p, C.c_int]),
    "cs_launch_count": (C.c_int64, [_VP]),
    "cs_step_kernel_trace": (C.c_int, [_VP, C.c_int, C.POINTER(C.c_uint64)]),
    "cs_stage_kernel_info": (C.c_int, [_VP, C.POINTER(C.c_int32), C.POINTER(C.c_int64)]),
    "cs_diagnostics_noalloc": (C.c_int, [C.c_int] * 6 + [C.c_double, C.c_double, _VP, _VP, _VP, _VP, _VP, C.c_int, _VP]),
}

_lib = None


def exported_symbols():
    """Names the header declares; tests check each one is exported."""
    return sorted(_PROTOTYPES)


def lib():
    """Load the shared library once.  Raises RuntimeError when it is not built."""
    global _lib
    if _lib is None:
        path = Path(os.environ.get("CUBESPHERE_B200_LIB", LIB_PATH))
        if not path.exists():
            raise RuntimeError(
                f"{path} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(nvcc, sm_100a).  There is no CPU or PyTorch fallback for this path.")
        handle = C.CDLL(str(path))
        for name, (res, args) in _PROTOTYPES.items():
            fn = getattr(handle, name)
            fn.restype, fn.argtypes = res, args
        if handle.cs_abi_version() != CS_ABI_VERSION:
            raise RuntimeError(f"{path}: ABI version {handle.cs_abi_version()} != {CS_ABI_VERSION}")
        _lib = handle
    return _lib


def check(rc: int, ctx=None):
    """Turn a negative return code into the Python exception the reference would raise."""
    if rc == CS_OK:
        return
    msg = lib().cs_last_error(ctx)
    msg = msg.decode() if msg else f"error {rc}"
    if rc == CS_E_INVALID:
        raise ValueError(msg)
    if rc == CS_E_UNSUPPORTED:
        raise NotImplementedError(msg)
    if rc == CS_E_NOMEM:
        raise MemoryError(msg)
    raise RuntimeError(msg)
