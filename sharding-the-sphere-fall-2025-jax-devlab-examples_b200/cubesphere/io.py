"""History / restart files for the solver state (host side).

The reference names zarr history files and Orbax restarts (deck slides 4 and 6) but ships neither, and
neither package exists in this image.  The state is three caller-owned arrays, so a restart is one
compressed ``.npz`` with the arrays and the metadata needed to rebuild the context; a history is a
directory of such files.  Arrays are stored in the padded tile layout ``[tiles, n + 2g, n + 2g]``.
"""
from __future__ import annotations

import json
from pathlib import Path

import numpy as np

FORMAT = "cubesphere-b200-state-v1"


def _to_numpy(a):
    if hasattr(a, "detach"):            # torch tensor (CPU or CUDA)
        a = a.detach().cpu().numpy()
    return np.ascontiguousarray(a)


def save_state(path, state, *, N, tiles_per_edge=1, halo=2, step=0, time=0.0, case=None, extra=None):
    """Write ``state = {'h', 'u', 'v'}`` (NumPy arrays or torch tensors) plus metadata to ``path`` (.npz)."""
    arrays = {k: _to_numpy(state[k]) for k in ("h", "u", "v")}
    shape = arrays["h"].shape
    n = N // tiles_per_edge
    if any(a.shape != shape for a in arrays.values()) or shape[1:] != (n + 2 * halo, n + 2 * halo):
        raise ValueError(f"state arrays must share the shape [tiles, {n + 2 * halo}, {n + 2 * halo}], got {shape}")
    meta = dict(format=FORMAT, N=int(N), tiles_per_edge=int(tiles_per_edge), halo=int(halo), step=int(step),
                time=float(time), case=case, dtype=str(arrays["h"].dtype), extra=extra or {})
    path = Path(path)
    path.parent.mkdir(parents=True, exist_ok=True)
    np.savez_compressed(path, meta=np.frombuffer(json.dumps(meta).encode(), dtype=np.uint8), **arrays)
    return path


def load_state(path):
    """Return ``(state, meta)``; ``state`` holds NumPy arrays in the padded tile layout."""
    with np.load(Path(path)) as f:
        meta = json.loads(bytes(f["meta"]).decode())
        if meta.get("format") != FORMAT:
            raise ValueError(f"{path}: not a {FORMAT} file")
        state = {k: f[k] for k in ("h", "u", "v")}
    return state, meta
