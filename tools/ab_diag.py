#!/usr/bin/env python
"""Tuning aid: time cs_diagnostics at C1536 for several builds of the library (one process each)."""
import os
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
if os.environ.get("AB_CHILD"):
    sys.path.insert(0, str(ROOT / "sharding-the-sphere-fall-2025-jax-devlab-examples_b200"))
    import torch
    from cubesphere import api
    res = []
    for dtype in ("f32", "f64"):
        ctx = api.Context(1536, halo=2, dtype=dtype)
        st = [torch.full(ctx.shape, v, dtype=ctx.dtype, device="cuda") for v in (1000.0, 1e-6, 2e-6)]
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for _ in range(3):
            d = ctx.diagnostics(*st)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(20):
            d = ctx.diagnostics(*st)
        e1.record()
        torch.cuda.synchronize()
        res.append(f"{dtype}: {e0.elapsed_time(e1) / 20 * 1e3:7.2f} us  mass {d[0].item():.15e}")
        ctx.close()
    print(" | ".join(res), flush=True)
else:
    for lib in sys.argv[1:]:
        env = dict(os.environ, AB_CHILD="1", CUBESPHERE_B200_LIB=str(Path(lib).resolve()))
        r = subprocess.run([sys.executable, __file__], env=env, capture_output=True, text=True)
        print(f"{Path(lib).name:>12}: {r.stdout.strip() or r.stderr[-300:]}", flush=True)
