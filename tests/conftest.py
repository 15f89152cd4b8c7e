import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "2022a-genome-dynamics_b200")
for p in (PKG, ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")
    config.addinivalue_line("markers", "slow: longer statistical checks")


@pytest.fixture(scope="session", autouse=True)
def _built_libraries():
    """Make sure the CPU-side checker libraries exist (cheap no-op when already built)."""
    import subprocess

    subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "oracle")], check=True,
                   stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    yield


@pytest.fixture(scope="session")
def oracle():
    import checkers

    return checkers.load_oracle()


@pytest.fixture(scope="session")
def reference():
    import checkers

    if not checkers.have_reference():
        pytest.skip("oracle/_ref/libmicromd_ref.so not built (no /root/reference here)")
    return checkers.load_reference()


@pytest.fixture(scope="session")
def engine():
    import bdb200
    import torch

    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return bdb200.load_engine()  # raises loudly if the CUDA library is missing
