import importlib
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def fdls():
    """The product package (directory name starts with a digit -> importlib)."""
    return importlib.import_module("4dlangsplat_b200")


@pytest.fixture(scope="session")
def synth():
    return importlib.import_module("4dlangsplat_b200.synth")
