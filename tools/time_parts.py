#!/usr/bin/env python
"""Timing aid: the two parts of a split stage (inner: marching kernel, rim: per-cell kernel) of one shard, with and
without the fused conservation sums.  One GPU; the shard's remote ghosts are simply left as they are."""
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "sharding-the-sphere-fall-2025-jax-devlab-examples_b200"))
import torch  # noqa: E402
from cubesphere import api  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 2176
dtype = sys.argv[2] if len(sys.argv) > 2 else "f32"
tiles = int(sys.argv[3]) if len(sys.argv) > 3 else 3
ctx = api.Context(N, halo=2, dtype=dtype, tile_begin=0, tile_count=tiles)
gen = torch.Generator(device="cuda").manual_seed(1)
q = [1000.0 + 10.0 * torch.rand(ctx.shape, device="cuda", dtype=ctx.dtype, generator=gen),
     1e-6 * torch.rand(ctx.shape, device="cuda", dtype=ctx.dtype, generator=gen),
     1e-6 * torch.rand(ctx.shape, device="cuda", dtype=ctx.dtype, generator=gen)]
q0 = [x.clone() for x in q]
out = [ctx.new_field() for _ in range(3)]
d = torch.zeros(2, dtype=torch.float64, device="cuda")
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)


def timed(fn, reps=10):
    fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(reps):
            fn()
    g.replay()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(3):
        e0.record()
        g.replay()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / reps)
    return best * 1e3


for part in (0, 1, 2):
    for a in (0.0, 0.75):
        t0 = timed(lambda: ctx.rk_stage(q0, q, out, a, 1 - a, 5.0, part=part))
        t1 = timed(lambda: ctx.rk_stage(q0, q, out, a, 1 - a, 5.0, part=part, diag=d, sums_of_input=True))
        t2 = timed(lambda: ctx.rk_stage(q0, q, out, a, 1 - a, 5.0, part=part, diag=d, sums_of_input=False))
        print(f"part {part} a={a}: plain {t0:7.1f} us | sums of input {t1:7.1f} | sums of output {t2:7.1f}")
ctx.close()
