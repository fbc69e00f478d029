import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
PKG = ROOT / "sharding-the-sphere-fall-2025-jax-devlab-examples_b200"
for p in (str(PKG), str(ROOT)):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def built_library():
    """The in-tree shared library; built on demand here (nvcc cross-compiles without a GPU)."""
    from cubesphere import _binding
    if not _binding.LIB_PATH.exists():
        import __graft_entry__ as entry
        entry.build()
    return _binding.lib()
