import importlib
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu)")


@pytest.fixture(scope="session")
def pkg():
    """The product package (directory name has a hyphen, so importlib)."""
    return importlib.import_module("vllm-quick_b200")


@pytest.fixture(scope="session")
def oracle():
    from oracle import oracle as o
    o.build()
    return o


@pytest.fixture(scope="session")
def cuda_lib(pkg):
    """Build (if needed) and load libb200_w4a16.so; never silently skip."""
    if not os.path.exists(pkg._lib.LIB_PATH):
        pkg._lib.build()
    return pkg._lib.lib()
