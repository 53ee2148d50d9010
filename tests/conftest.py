import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session", autouse=True)
def _built():
    """Make sure the oracle libraries and the CUDA library exist (no-ops when up to date)."""
    oracle_so = os.path.join(ROOT, "oracle", "_ref", "libnsdg_oracle.so")
    gpu_so = os.path.join(ROOT, "nextsimdg_b200", "csrc", "libnsdg_gpu.so")
    if not (os.path.exists(oracle_so) and os.path.exists(gpu_so)):
        import __graft_entry__ as g
        g.build()
    yield


def has_cuda():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.fixture(scope="session")
def ctx():
    from nextsimdg_b200.gpu import Context
    c = Context(0)
    yield c
    c.close()
