"""Registration of the C ABI as ``jax.ffi`` custom-call targets -- the drop-in the north star names.

Usable only where JAX and the XLA-FFI adapter (``csrc/xla_ffi_shim.cc`` -> ``lib/libcubesphere_xla.so``)
exist; neither does in this image, so nothing here is imported by the package or covered by the tests.
It mirrors what :mod:`cubesphere.api` does with torch tensors: same names as the reference
(``make_halo_exchange`` JAX-DevLab-Examples.py:199-246; ``step(state, dt)`` of BASELINE.json).
"""
from __future__ import annotations

import ctypes
from pathlib import Path

from . import _binding as B


def register():
    """Load the adapter and register its handlers.  Raises ImportError without JAX, RuntimeError without the adapter."""
    import jax  # noqa: F401  (ImportError here is the honest answer on a box without JAX)
    import jax.ffi

    path = B.PKG_ROOT / "lib" / "libcubesphere_xla.so"
    if not Path(path).exists():
        raise RuntimeError(f"{path} is missing: build it with `make -C csrc shim XLA_FFI_INCLUDE=...`")
    shim = ctypes.CDLL(str(path))
    for name, sym in (("cs_step", "CsStep"), ("cs_halo_fill_state", "CsHaloFillState"),
                      ("cs_halo_fill_scalar", "CsHaloFillScalar"), ("cs_diagnostics", "CsDiagnostics")):
        jax.ffi.register_ffi_target(name, jax.ffi.pycapsule(getattr(shim, sym)), platform="CUDA")
    return shim


def make_step(ctx_handle: int):
    """``step(state, dt) -> state`` on JAX arrays; ``ctx_handle`` is the integer value of a ``cs_ctx*``
    created through :class:`cubesphere.api.Context` (``ctx._h.value``) for the device the arrays live on."""
    import jax
    import jax.ffi

    def step(state, dt):
        ops = (state["h"], state["u"], state["v"])
        shapes = tuple(jax.ShapeDtypeStruct(x.shape, x.dtype) for x in ops)
        h, u, v = jax.ffi.ffi_call("cs_step", shapes, input_output_aliases={0: 0, 1: 1, 2: 2})(
            *ops, ctx=int(ctx_handle), dt=float(dt))
        return {"h": h, "u": u, "v": v}

    return step


def make_halo_exchange(ctx_handle: int):
    """``field_ghosts -> field_ghosts`` as one custom call (the reference's compiled closure, PY:238-246)."""
    import jax
    import jax.ffi

    def cubesphere_halo_exchange(field_ghosts):
        return jax.ffi.ffi_call("cs_halo_fill_scalar", jax.ShapeDtypeStruct(field_ghosts.shape, field_ghosts.dtype),
                                input_output_aliases={0: 0})(field_ghosts, ctx=int(ctx_handle))

    return cubesphere_halo_exchange
