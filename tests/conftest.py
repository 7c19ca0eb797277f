import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def oracle():
    import oracle_lib
    oracle_lib.lib()
    return oracle_lib


@pytest.fixture(scope="session")
def pkg():
    import mcb200_loader
    return mcb200_loader.load()


@pytest.fixture(scope="session")
def gpu(pkg):
    """the initialised CUDA engine; fails loudly (never falls back) when the library or the GPU is missing"""
    pkg.engine.init()
    yield pkg
    pkg.engine.shutdown()
