import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

import tsb200_loader  # noqa: E402


def _build_missing_native_pieces():
    """The in-tree libraries are build products (git-ignored): when a checkout has none yet -- the suite was
    started before __graft_entry__.build() -- make them here, before the package is imported.  On the GPU box
    they travel with the snapshot and nothing is built."""
    import shutil
    import subprocess

    csrc = os.path.join(ROOT, "torchsnapshot-1_b200", "csrc")
    missing = [t for t in ("libtsb200.so", "_tsb200_meta.so") if not os.path.exists(os.path.join(csrc, t))]
    if missing and shutil.which("make") and (shutil.which("nvcc") or os.path.exists("/usr/local/cuda/bin/nvcc")):
        subprocess.run(["make", "-C", csrc] + missing, check=False, capture_output=True)


_build_missing_native_pieces()
tsb200_loader.load()


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on the B200 box)")


@pytest.fixture(scope="session")
def golden_dir():
    return os.path.join(ROOT, "tests", "golden")


@pytest.fixture(scope="session", autouse=True)
def _oracle_built():
    import oracle

    oracle.lib()
    yield


@pytest.fixture
def engine():
    # function scope on purpose: a test may shut the engines down (to change arena bounds), and
    # get_engine() hands back the cached engine or builds a new one
    import torch

    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from torchsnapshot_b200.engine import get_engine

    return get_engine(0)
