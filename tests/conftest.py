import importlib
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def pkg():
    """The product package (hyphenated directory name)."""
    return importlib.import_module("gpu-radix-sort_b200")


@pytest.fixture(scope="session")
def oracle():
    from oracle import oracle as o
    o.lib()
    return o


@pytest.fixture(scope="session")
def libsort(pkg):
    """libsort.so, built if needed.  No skip: a missing library is a failure."""
    if not os.path.exists(pkg.lib_path()):
        pkg.build()
    return pkg.lib()


@pytest.fixture(scope="session")
def gpu_lib(libsort, pkg):
    pkg.binding.init_once()
    return libsort
