import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "sparse-solvers_b200"))
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN


def have_ref():
    from oracle import ref
    return ref.available() and os.path.exists("/root/repo/oracle/_ref/libcblas_fwd.so")


needs_ref = pytest.mark.skipif(not have_ref(), reason="oracle/_ref not built (needs /root/reference at build time)")
