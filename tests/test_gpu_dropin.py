"""The reference's own pybind11 module (binding.cpp, unchanged) on top of libssb200 -- the drop-in check."""
import os
import subprocess
import sys

import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu
BUILT = os.path.join(ROOT, "dropin", "_build", "sparsesolvers")


@pytest.mark.skipif(not os.path.isdir(BUILT), reason="dropin/_build not built (needs /root/reference headers at build time)")
def test_reference_binding_tests_pass_on_the_drop_in():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "dropin", "run_binding_tests.py")], capture_output=True, text=True,
                       timeout=600)
    print(r.stdout[-2000:], r.stderr[-3000:])
    assert r.returncode == 0
