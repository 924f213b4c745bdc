import importlib
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session", autouse=True)
def _built():
    """Build the native artefacts if a fresh checkout has none (they are git-ignored; `.so` files travel to the GPU box)."""
    import __graft_entry__ as g
    wanted = [g.LIB, g.HOST_LIB] + [os.path.join(g.BIN, exe) for exe in ("igtrace", "igcli", "igdbtool")]
    if not all(os.path.exists(p) for p in wanted):
        g.build()


@pytest.fixture(scope="session")
def host_bin():
    import __graft_entry__ as g
    return g.BIN


@pytest.fixture(scope="session")
def ig():
    """The product package (directory name has a hyphen, hence importlib)."""
    return importlib.import_module("ignis-masterthesis_b200")


@pytest.fixture(scope="session")
def oracle():
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import pyoracle
    pyoracle.load()
    return pyoracle
