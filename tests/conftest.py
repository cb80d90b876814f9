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
def ref_fixtures():
    """(module, arrays): outputs of the reference's own CPU kernels on seeded inputs (tests/golden/ref_outputs.npz,
    written by tests/golden/make_ref_fixtures.py where /root/reference is mounted) and the module that regenerates
    the inputs."""
    import importlib.util
    import numpy as np
    path = os.path.join(ROOT, "tests", "golden", "make_ref_fixtures.py")
    spec = importlib.util.spec_from_file_location("make_ref_fixtures", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod, np.load(os.path.join(ROOT, "tests", "golden", "ref_outputs.npz"))


@pytest.fixture(scope="session")
def pkg():
    """The product package (directory name has a hyphen, so it is loaded by path)."""
    import __graft_entry__ as ge
    return ge.load_package()
