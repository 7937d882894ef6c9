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
def sba():
    """The product package (hyphenated directory name -> importlib)."""
    return importlib.import_module("semantic-bundle-adjustment-colmap_b200")


@pytest.fixture(scope="session")
def oracle(sba):
    """CPU restatement (checker). Built on demand with oracle/Makefile."""
    from oracle import loader
    return loader.load(sba.capi)
