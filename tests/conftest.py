import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def q():
    """The ctypes plumbing module of the product library (hpg-fastq_b200/hpgq.py)."""
    import __graft_entry__ as g

    so = os.path.join(g.PKG_DIR, "libhpgq_cuda.so")
    if not os.path.exists(so):
        g.build()
    return g.load_package().hpgq


@pytest.fixture(scope="session")
def ob():
    import oracle_binding

    oracle_binding.lib()
    return oracle_binding
