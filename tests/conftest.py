import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box: pytest -m gpu)")


@pytest.fixture(scope="session")
def oracle_lib():
    from tests import orc
    return orc.lib()


@pytest.fixture(scope="session")
def gpu_ctx_factory():
    """Context factory; fails loudly if the CUDA extension is missing (no CPU fallback)."""
    import liverrenderer_b200 as mi

    def make(scene):
        return mi.Context(0).upload(scene)
    return make
