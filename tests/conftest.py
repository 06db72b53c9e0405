import importlib
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

PKG = "distributed-real-time-chat-and-collaboration-tool_b200"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


def pkg(sub: str | None = None):
    """Import the product package (its directory name has hyphens) or one of its submodules."""
    return importlib.import_module(PKG if sub is None else f"{PKG}.{sub}")


@pytest.fixture(scope="session")
def ops():
    return pkg("ops")


@pytest.fixture(scope="session")
def cuda():
    import torch

    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    torch.cuda.set_device(0)
    return torch.device("cuda:0")
