#!/usr/bin/env python
"""Tuning aid: time the fused stage kernel of several builds of the library on the same GPU, interleaved.
usage: ab_stage.py lib_a.so lib_b.so ...   (each measured in its own process, three rounds)"""
import os
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent


def child():
    sys.path.insert(0, str(ROOT / "sharding-the-sphere-fall-2025-jax-devlab-examples_b200"))
    import torch
    from cubesphere import api
    N = int(os.environ.get("AB_N", "1536"))
    out_line = []
    for dtype in ("f32", "f64"):
        ctx = api.Context(N, halo=2, dtype=dtype)
        gen = torch.Generator(device="cuda").manual_seed(1)
        h = 1000.0 + 10.0 * torch.rand(ctx.shape, device="cuda", dtype=ctx.dtype, generator=gen)
        u = 1e-6 * torch.rand(ctx.shape, device="cuda", dtype=ctx.dtype, generator=gen)
        v = 1e-6 * torch.rand(ctx.shape, device="cuda", dtype=ctx.dtype, generator=gen)
        ctx.halo_fill_state(h, u, v)
        out = [ctx.new_field() for _ in range(3)]
        q0 = [x.clone() for x in (h, u, v)]
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for a, b in ((0.0, 1.0), (0.75, 0.25)):
            best = 1e9
            for _ in range(3):
                for _ in range(3):
                    ctx.rk_stage(q0, (h, u, v), out, a, b, 10.0, part=2)
                torch.cuda.synchronize()
                e0.record()
                for _ in range(20):
                    ctx.rk_stage(q0, (h, u, v), out, a, b, 10.0, part=2)
                e1.record()
                torch.cuda.synchronize()
                best = min(best, e0.elapsed_time(e1) / 20)
            out_line.append(f"{dtype} a={a:g}: {best * 1e3:7.2f} us")
        ctx.close()
    print(" | ".join(out_line), flush=True)


if __name__ == "__main__":
    if os.environ.get("AB_CHILD"):
        child()
    else:
        libs = sys.argv[1:]
        for rnd in range(int(os.environ.get("AB_ROUNDS", "2"))):
            for lib in libs:
                env = dict(os.environ, AB_CHILD="1", CUBESPHERE_B200_LIB=str(Path(lib).resolve()))
                r = subprocess.run([sys.executable, __file__], env=env, capture_output=True, text=True)
                print(f"{Path(lib).name:>16}: {r.stdout.strip() or r.stderr[-400:]}", flush=True)
