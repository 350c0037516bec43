"""pytest wiring: `gpu` marker, import helpers, shared fixtures."""
import importlib.util
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def load_pkg():
    """import llama.cpp_test_b200/ (dotted directory name) as module `llama_cpp_test_b200`"""
    name = "llama_cpp_test_b200"
    if name in sys.modules:
        return sys.modules[name]
    path = os.path.join(ROOT, "llama.cpp_test_b200")
    spec = importlib.util.spec_from_file_location(name, os.path.join(path, "__init__.py"),
                                                  submodule_search_locations=[path])
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


@pytest.fixture(scope="session")
def pkg():
    return load_pkg()


@pytest.fixture(scope="session")
def oracle():
    from oracle.loader import Oracle
    return Oracle()


@pytest.fixture(scope="session")
def ref():
    from oracle.loader import Ref
    r = Ref.load()
    if r is None:
        pytest.skip("oracle/_ref not built (reference tree absent when the checker was built)")
    return r


@pytest.fixture(scope="session")
def cuda(pkg):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch.device("cuda:0")
