import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    config.addinivalue_line("markers", "slow: takes more than a few seconds")


@pytest.fixture(scope="session", autouse=True)
def _build_oracle():
    """The oracle is test infrastructure: (re)build its C restatement when missing."""
    odir = os.path.join(ROOT, "oracle")
    if not os.path.exists(os.path.join(odir, "libparsim_oracle.so")):
        subprocess.run(["make", "-C", odir], check=True, capture_output=True)


def has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False
