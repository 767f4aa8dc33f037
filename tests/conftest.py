import os
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
for p in (str(ROOT), str(ROOT / "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu on the GPU box")


@pytest.fixture(scope="session")
def lib():
    """The C-ABI library, built in-tree if missing (nvcc cross-compiles without a GPU)."""
    from multimodal_diffusion_b200 import _lib, build
    if not _lib.LIB_PATH.exists():
        build.build()
    return _lib.load()
