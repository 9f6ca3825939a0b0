import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session", autouse=True)
def _built_libraries():
    """nvcc / gcc are available without a GPU: make sure the product library and
    the oracle exist before any test imports them."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("_guava_build", os.path.join(ROOT, "guava_renderer_b200", "build.py"))
    gb = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(gb)      # not `import guava_renderer_b200.build`: the package refuses to
    gb.build()                       # import without its CUDA library (no fallback)
    import oracle
    oracle.build()
    yield


def has_cuda():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False
