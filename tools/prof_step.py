#!/usr/bin/env python
"""Profiling aid: a few whole time steps at a given size / dtype through cs_step (run under ncu).
CS_STEP_KERNEL=0 in the environment selects the one-stage-per-launch path."""
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "sharding-the-sphere-fall-2025-jax-devlab-examples_b200"))
import os  # noqa: E402
import torch  # noqa: E402
from cubesphere import api  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 1536
dtype = sys.argv[2] if len(sys.argv) > 2 else "f32"
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
ctx = api.Context(N, halo=2, dtype=dtype, sums_of_entry_state=True, fused_halo=bool(os.environ.get("PROF_FUSED_HALO")))
gen = torch.Generator(device="cuda").manual_seed(1)
h = 1000.0 + 10.0 * torch.rand(ctx.shape, device="cuda", dtype=ctx.dtype, generator=gen)
u = 1e-6 * torch.rand(ctx.shape, device="cuda", dtype=ctx.dtype, generator=gen)
v = 1e-6 * torch.rand(ctx.shape, device="cuda", dtype=ctx.dtype, generator=gen)
ctx.halo_fill_state(h, u, v)
import os
diag = None if os.environ.get("PROF_NODIAG") else torch.zeros(2, dtype=torch.float64, device="cuda")   # PROF_NODIAG=1: no sums
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
ctx.step(h, u, v, 5.0, diag=diag)
torch.cuda.synchronize()
e0.record()
for _ in range(steps):
    ctx.step(h, u, v, 5.0, diag=diag)
e1.record()
torch.cuda.synchronize()
import os
if int(os.environ.get("CS_STEP_DEBUG", "0")) & 16:
    import ctypes as C
    import numpy as np
    nb = 293 if len(sys.argv) < 5 else int(sys.argv[4])
    tr = np.zeros((nb, 8), dtype=np.uint64)
    ctx._check(ctx._lib.cs_step_kernel_trace(ctx._h, nb, tr.ctypes.data_as(C.POINTER(C.c_uint64))))
    t = (tr.astype(np.int64) - int(tr[:, 0].min())) / 1e3
    names = ["start", "march1", "bar1", "march2", "bar2", "march3", "bar3", "fill"]
    d1 = t[:, 1] - t[:, 0]
    order = np.argsort(d1)
    print("  march1 by block range:", [round(float(d1[a:a + 37].mean()), 1) for a in range(0, nb, 37)])
    print("  slowest blocks", order[-12:].tolist(), "fastest", order[:12].tolist())
    for k, nm in enumerate(names):
        print(f"  {nm:7s} min {t[:, k].min():8.1f}  mean {t[:, k].mean():8.1f}  max {t[:, k].max():8.1f} us")
print("ok", diag.tolist() if diag is not None else None, "ms/step", e0.elapsed_time(e1) / steps, "launches", ctx.launches)
ctx.close()
