import importlib.util
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def load_package():
    """import the product package (directory name is not an identifier) as `eagle_b200`"""
    if "eagle_b200" in sys.modules:
        return sys.modules["eagle_b200"]
    path = os.path.join(ROOT, "eagle-in-llama.cpp_b200", "__init__.py")
    spec = importlib.util.spec_from_file_location("eagle_b200", path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules["eagle_b200"] = mod
    spec.loader.exec_module(mod)
    return mod


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def eb():
    mod = load_package()
    mod.init(0)          # raises without an sm_100 device: no CPU fallback
    return mod


@pytest.fixture(scope="session")
def po():
    import pyoracle
    pyoracle.oracle()
    return pyoracle
