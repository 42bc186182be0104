import os
import sys

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
for p in (ROOT, HERE):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on the B200 box)")


@pytest.fixture(scope="session")
def built_library():
    """The CUDA library must exist (build() makes it); tests never fall back to anything else."""
    from sparse_conv_gpis_tungsten_b200 import abi, build
    if not os.path.exists(abi.LIB_PATH):
        build.build_library()
    return abi.load()
