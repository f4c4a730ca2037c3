import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (sm_100a) GPU; run with -m gpu on the GPU box")


@pytest.fixture(scope="session")
def golden():
    from safetensors.torch import load_file

    def _load(name):
        return load_file(os.path.join(GOLDEN, name + ".safetensors"))
    return _load


@pytest.fixture(scope="session", autouse=True)
def _built_library():
    """the C-ABI library must exist for every tier (it is built by __graft_entry__.build())"""
    from hivae_b200 import _lib
    if not os.path.exists(_lib.LIB_PATH):
        from hivae_b200 import build
        build.build()
