#!/usr/bin/env python
"""Tuning aid: whole-step time (cs_step, graph replay) of several builds of the library on the same GPU, interleaved.
usage: ab_step.py lib_a.so lib_b.so ..."""
import os
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
if __name__ == "__main__":
    libs = sys.argv[1:]
    for rnd in range(int(os.environ.get("AB_ROUNDS", "2"))):
        for lib in libs:
            for dtype in ("f32", "f64"):
                env = dict(os.environ, CUBESPHERE_B200_LIB=str(Path(lib).resolve()))
                r = subprocess.run([sys.executable, str(ROOT / "tools" / "prof_step.py"), os.environ.get("AB_N", "1536"), dtype, "100"],
                                   env=env, capture_output=True, text=True)
                out = [l for l in r.stdout.splitlines() if "ms/step" in l]
                print(f"{Path(lib).name:>16} {dtype}: {out[-1][out[-1].index('ms/step'):] if out else r.stderr[-300:]}", flush=True)
