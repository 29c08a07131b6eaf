import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def oracle():
    import oracle as O
    O.build()
    return O


@pytest.fixture(scope="session")
def rt():
    """The product package; the GPU tests call through its ctypes C ABI."""
    import gpu_raytracer_b200 as g
    g.lib()
    return g


def bits(a):
    """uint32 view for bit-exact comparisons of float arrays."""
    return np.ascontiguousarray(a).view(np.uint32)
