"""pytest configuration: registers the `gpu` marker and puts the product directory on sys.path.

`-m "not gpu"` runs everywhere (oracle vs golden fixtures, host logic, C-ABI symbol checks);
`-m gpu` needs a B200 and calls the CUDA path through the C-ABI.
"""
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "fp8-emulation-toolkit_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_collection_modifyitems(config, items):
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:  # pragma: no cover
        has_gpu = False
    if has_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


def pytest_sessionstart(session):
    """Build the native pieces when they are missing (fresh checkout): libfpemu_b200.so needs nvcc (cross-compiles
    without a GPU), the C oracle needs gcc.  Nothing is rebuilt when the files are already there."""
    import shutil
    lib = os.path.join(PKG, "libfpemu_b200.so")
    if not os.path.exists(lib) and (shutil.which("nvcc") or os.path.exists("/usr/local/cuda/bin/nvcc")):
        import importlib.util
        spec = importlib.util.spec_from_file_location("fpemu_b200_build", os.path.join(PKG, "build.py"))
        b = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(b)
        b.build()
    try:
        from oracle import fpemu_oracle
        fpemu_oracle.build()
    except Exception:   # pragma: no cover  (reported by the tests that need it)
        pass


@pytest.fixture(scope="session")
def golden_dir():
    return os.path.join(ROOT, "tests", "golden")
