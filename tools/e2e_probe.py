#!/usr/bin/env python
"""Tuning aid: end-to-end time per member of cs_step_host vs cs_step_host_batch (same / distinct pinned buffers)."""
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "sharding-the-sphere-fall-2025-jax-devlab-examples_b200"))
import torch  # noqa: E402
from cubesphere import api  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 1536
ctx = api.Context(N, halo=2, dtype="f32")
mk = lambda s: [torch.full(ctx.shape, v, dtype=ctx.dtype).pin_memory() for v in (1000.0 + s, 1e-6, 1e-6)]
ens = [mk(i) for i in range(4)]
dt = 5.0


def wall(fn, n):
    fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) * 1e3 / n


NS = int(sys.argv[2]) if len(sys.argv) > 2 else 1
for cb in (False, True):
    print("copy_back", cb)
    print("  single x8         %.3f ms" % wall(lambda: [ctx.step_host(*ens[0], dt, nsteps=NS, copy_back=cb) for _ in range(8)], 8))
    print("  batch same buffer %.3f ms" % wall(lambda: ctx.step_host_batch([ens[0]] * 8, dt, nsteps=NS, copy_back=cb), 8))
    print("  batch 4 distinct  %.3f ms" % wall(lambda: ctx.step_host_batch(ens * 2, dt, nsteps=NS, copy_back=cb), 8))
    print("  batch 2 distinct  %.3f ms" % wall(lambda: ctx.step_host_batch(ens[:2] * 4, dt, nsteps=NS, copy_back=cb), 8))
# raw copy speed of each buffer
dev = torch.empty(ctx.shape, dtype=ctx.dtype, device="cuda")
for i, m in enumerate(ens):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    dev.copy_(m[0], non_blocking=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(5):
        dev.copy_(m[0], non_blocking=True)
    e1.record()
    torch.cuda.synchronize()
    print("  H2D member %d: %.1f GB/s" % (i, 5 * m[0].numel() * 4 / (e0.elapsed_time(e1) * 1e6)))
ctx.close()
