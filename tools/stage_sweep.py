#!/usr/bin/env python
"""Tuning aid: time the fused stage kernel alone for combinations of the marching kernel's
environment knobs (CS_MARCH_VEC, CS_MARCH_ROWS).  Not part of the product or the tests."""
import itertools
import os
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "sharding-the-sphere-fall-2025-jax-devlab-examples_b200"))
import torch  # noqa: E402
from cubesphere import api  # noqa: E402


def main():
    N = int(sys.argv[1]) if len(sys.argv) > 1 else 1536
    for dtype, esize in (("f32", 4), ("f64", 8)):
        for kernel in ("march",):
            ctx = api.Context(N, halo=2, dtype=dtype, kernel=kernel)
            gen = torch.Generator(device="cuda").manual_seed(1)
            h = 1000.0 + 10.0 * torch.rand(ctx.shape, device="cuda", dtype=ctx.dtype, generator=gen)
            u = 1e-6 * torch.rand(ctx.shape, device="cuda", dtype=ctx.dtype, generator=gen)
            v = 1e-6 * torch.rand(ctx.shape, device="cuda", dtype=ctx.dtype, generator=gen)
            ctx.halo_fill_state(h, u, v)
            out = [ctx.new_field() for _ in range(3)]
            q0 = [x.clone() for x in (h, u, v)]          # distinct arrays: in a real step q0 != qs for stages 2, 3
            if dtype == "f32":
                continue
            combos = [(1, 0, 0, 2), (1, 32, 0, 2), (1, 64, 0, 2), (2, 0, 2, 2)]
            for vec, rows, occ, ring in combos:
                if True:
                    os.environ["CS_MARCH_VEC"], os.environ["CS_MARCH_OCC"], os.environ["CS_TMA_RING"] = str(vec or 4), str(occ), str(ring or 6)
                    if rows:
                        os.environ["CS_MARCH_ROWS"] = str(rows)
                    else:
                        os.environ.pop("CS_MARCH_ROWS", None)
                    os.environ["CS_MARCH_PREFETCH"] = str(ring)
                res = []
                for a, words in ((0.0, 6), (0.75, 9)):
                    for _ in range(2):
                        ctx.rk_stage(q0, (h, u, v), out, a, 1 - a, 5.0, part=2)
                    torch.cuda.synchronize()
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record()
                    for _ in range(10):
                        ctx.rk_stage(q0, (h, u, v), out, a, 1 - a, 5.0, part=2)
                    e1.record()
                    torch.cuda.synchronize()
                    ms = e0.elapsed_time(e1) / 10
                    res.append((ms, words * esize * 6 * N * N / ms / 1e6))
                print(f"{dtype} {kernel} vec={vec} occ={occ} rows={rows} ring/pf={ring}: stage1 {res[0][0]:.4f} ms {res[0][1]:.0f} GB/s | stage2 {res[1][0]:.4f} ms {res[1][1]:.0f} GB/s", flush=True)
            ctx.close()


if __name__ == "__main__":
    main()
