import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

import __graft_entry__ as graft  # noqa: E402

# torch first, where it is installed: its bundled NCCL (2.28) must be the libnccl.so.2 of the process before csrc/comm.cu
# dlopen()s one — otherwise, on a multi-GPU box, the system's older NCCL gets loaded by the film-reduce tests and a later
# `import torch` in the same process fails on a symbol it lacks (seen on the 2-GPU box: ncclDevCommCreate)
try:
    import torch  # noqa: F401,E402
except ImportError:
    pass


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def pkg():
    return graft.load_package()


@pytest.fixture(scope="session")
def ctx(pkg):
    """One GPU context for the whole session; fails loudly (no skip, no fallback) if there is no device."""
    c = pkg.binding.Context(0)
    yield c
    c.close()


@pytest.fixture(scope="session")
def golden():
    def load(name):
        return np.load(os.path.join(ROOT, "tests", "golden", name))
    return load
