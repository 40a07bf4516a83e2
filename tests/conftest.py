import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_collection_modifyitems(config, items):
    # files named test_zz_* always run last, whatever order pytest was given them in
    items.sort(key=lambda it: os.path.basename(str(it.fspath)).startswith("test_zz_"))


@pytest.fixture(scope="session")
def oracle():
    from oracle import oracle as O

    O.build()
    return O


def _product():
    import importlib

    return importlib.import_module("triton-kubernetes_b200")


@pytest.fixture(scope="session")
def tk():
    """The product package (C ABI via ctypes).  Importing it needs the built library, not a GPU."""
    return _product()


@pytest.fixture(params=["oracle", pytest.param("gpu", marks=pytest.mark.gpu)])
def impl(request, oracle):
    """Runs a test once against the oracle (CPU) and once against the product on the GPU."""
    return oracle if request.param == "oracle" else _product()
