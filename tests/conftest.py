import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def built_lib():
    """The in-tree CUDA library, (re)built if stale.  Building needs nvcc only, no GPU."""
    from oap_mllib_b200 import build
    path = build.build()
    assert os.path.exists(path)
    return path


@pytest.fixture(scope="session")
def oracle_c():
    """The plain-C oracle, built from oracle/kmeans_oracle.c."""
    import ctypes
    import subprocess
    subprocess.run(["make", "-C", os.path.join(ROOT, "oracle")], check=True, capture_output=True)
    return ctypes.CDLL(os.path.join(ROOT, "oracle", "libkmeans_oracle.so"))
