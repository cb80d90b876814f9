import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def goldens():
    import json
    with open(os.path.join(ROOT, "tests", "golden", "goldens.json")) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def pkg():
    """The product package (directory name has a hyphen, so it is loaded by path)."""
    import __graft_entry__ as ge
    return ge.load_package()
