import importlib.util
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
PKG = os.path.join(ROOT, "llama.cpp-q_4_0_b200")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu)")


def load_binding():
    """The ctypes binding lives in a directory whose name is not a Python identifier."""
    name = "ggml_b200"
    if name in sys.modules:
        return sys.modules[name]
    spec = importlib.util.spec_from_file_location(name, os.path.join(PKG, "ggml_b200.py"))
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


@pytest.fixture(scope="session")
def golden():
    import numpy as np
    return np.load(os.path.join(ROOT, "tests", "golden", "q4_0_golden.npz"))


@pytest.fixture(scope="session")
def port():
    import oracle
    return oracle.port()


@pytest.fixture(scope="session")
def ref():
    import oracle
    return oracle.ref()


@pytest.fixture(scope="session")
def b200():
    """Initialised library on cuda:0.  Fails loudly (no skip, no fallback) when a -m gpu test runs
    without the extension or without a GPU."""
    mod = load_binding()
    mod.init(0)
    yield mod
    mod.shutdown()
