import os
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session", autouse=True)
def _built():
    """Make sure the in-tree libraries exist (the driver normally ran build() already)."""
    from montecarlo_pathtracing_b200 import build
    lib = ROOT / "montecarlo_pathtracing_b200" / "lib"
    have_nvcc = os.path.exists("/usr/local/cuda/bin/nvcc")
    if have_nvcc:
        build.build_all()
    else:
        assert (lib / "libmcpt.so").exists(), "libmcpt.so missing and no nvcc to build it"
    yield
