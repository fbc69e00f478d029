"""Multi-GPU parity under real NCCL: one process per GPU (torchrun), the sharded stepper -- rim / inner split, halo
exchange overlapped with the inner cells, the step replayed as a CUDA graph, conservation sums fused into the last
stage and combined by an all-reduce -- against the CPU oracle.  Skipped on a box with fewer than two GPUs (the
single-GPU sharding tests in test_gpu_parity.py cover the plan and the kernels with a copy in place of NCCL)."""
import os
import subprocess
import sys
from pathlib import Path

import pytest

torch = pytest.importorskip("torch")

pytestmark = pytest.mark.gpu

ROOT = Path(__file__).resolve().parent.parent


def _world_sizes():
    n = torch.cuda.device_count() if torch.cuda.is_available() else 0
    return [w for w in (2, 4, 8) if w <= n]


@pytest.mark.parametrize("world", [2, 4, 8])
def test_sharded_stepper_under_nccl(built_library, world):
    if world not in _world_sizes():
        pytest.skip(f"needs {world} visible GPUs")
    port = 29500 + world + (os.getpid() % 200)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(port), str(ROOT / "tests" / "multi_gpu_check.py")]
    res = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=900)
    sys.stdout.write(res.stdout[-4000:])
    assert res.returncode == 0, res.stdout[-4000:] + res.stderr[-4000:]
    assert "FAIL" not in res.stdout and res.stdout.count("-> ok") >= world
