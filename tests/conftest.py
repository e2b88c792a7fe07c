import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


def _have_gpu():
    try:
        import torch

        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _have_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def built_lib():
    """The in-tree libb200r.so; built on demand so `pytest` works from a clean checkout."""
    from generativeaiexamples_b200 import _lib

    if not os.path.exists(_lib.LIB_PATH):
        import __graft_entry__

        __graft_entry__.build()
    return _lib.load()


@pytest.fixture(autouse=True)
def _fresh_engine_switches():
    """libb200r reads its B200R_* environment switches once per process: tests that change them call
    ``_lib.reload_env()``; this puts the library back in step with the restored environment."""
    yield
    from generativeaiexamples_b200 import _lib

    if _lib._lib is not None:
        _lib.reload_env()
