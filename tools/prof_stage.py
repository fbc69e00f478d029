#!/usr/bin/env python
"""Profiling aid: a handful of fused-stage launches at a given size / dtype (run under ncu)."""
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "sharding-the-sphere-fall-2025-jax-devlab-examples_b200"))
import torch  # noqa: E402
from cubesphere import api  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 1536
dtype = sys.argv[2] if len(sys.argv) > 2 else "f32"
kernel = sys.argv[3] if len(sys.argv) > 3 else "auto"
ctx = api.Context(N, halo=2, dtype=dtype, kernel=kernel)
gen = torch.Generator(device="cuda").manual_seed(1)
h = 1000.0 + 10.0 * torch.rand(ctx.shape, device="cuda", dtype=ctx.dtype, generator=gen)
u = 1e-6 * torch.rand(ctx.shape, device="cuda", dtype=ctx.dtype, generator=gen)
v = 1e-6 * torch.rand(ctx.shape, device="cuda", dtype=ctx.dtype, generator=gen)
ctx.halo_fill_state(h, u, v)
out = [ctx.new_field() for _ in range(3)]
q0 = [x.clone() for x in (h, u, v)]
for a in (0.0, 0.75, 0.75, 0.0, 0.75):
    ctx.rk_stage(q0, (h, u, v), out, a, 1 - a, 5.0, part=2)
torch.cuda.synchronize()
print("ok", float(out[0].sum()))
ctx.close()
