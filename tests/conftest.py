"""pytest configuration: the `gpu` marker and loaders for the package (its directory name is not an identifier)."""
import importlib.util
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))
PKG_DIR = ROOT / "llama.cpp-s390x_b200"


def load_package():
    name = "llama_cpp_s390x_b200"
    if name in sys.modules:
        return sys.modules[name]
    spec = importlib.util.spec_from_file_location(name, PKG_DIR / "__init__.py", submodule_search_locations=[str(PKG_DIR)])
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def pkg():
    return load_package()


@pytest.fixture(scope="session")
def orc():
    from oracle import refenv
    return refenv.load_oracle()


@pytest.fixture(scope="session")
def ref():
    """the unmodified reference (oracle/_ref) or None when it did not travel"""
    from oracle import refenv
    return refenv.load_ref()


@pytest.fixture(scope="session")
def gpu_ctx(pkg):
    import torch
    if not torch.cuda.is_available():
        pytest.fail("-m gpu tests need a CUDA device; there is no fallback path")
    torch.cuda.init()
    ctx = pkg.Context(0, stream="torch")
    yield ctx
    ctx.close()
