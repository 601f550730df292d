import importlib.util
import os
import sys
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "colmap-pcd_b200")


def _load(name, path):
    if name in sys.modules:
        return sys.modules[name]
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu)")


@pytest.fixture(scope="session")
def orc():
    """CPU oracle (test infrastructure only)."""
    return _load("orc", os.path.join(ROOT, "oracle", "orc.py"))


@pytest.fixture(scope="session")
def synth():
    return _load("pcd_synth", os.path.join(PKG, "synth.py"))


@pytest.fixture(scope="session")
def pcd():
    """ctypes binding over the C ABI (product library)."""
    return _load("pcdb200", os.path.join(PKG, "pcdb200.py"))


@pytest.fixture(scope="session")
def ctx(pcd):
    c = pcd.Context(0)
    yield c
    c.close()
