"""CPU tests of the drop-in boundary: the C-ABI library loads, exports every symbol include/ssb200.h
declares, and fails loudly (never falls back) when no sm_100 device is present."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from conftest import ROOT


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "ssb200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(ssb200_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_the_reference_boundary():
    syms = _declared_symbols()
    for need in ("ssb200_homotopy_create_f32", "ssb200_homotopy_solve_f64", "ssb200_homotopy_solve_batch_f32",
                 "ssb200_irls_create_f64", "ssb200_irls_solve_batch_f32", "ssb200_destroy", "ssb200_last_error",
                 "ssb200_group_create_f32", "ssb200_group_solve_batch_f32", "ssb200_group_solve_batch_dev_f64", "ssb200_group_destroy",
                 "ssb200_homotopy_set_residual_mode", "ssb200_homotopy_get_residual_info"):
        assert need in syms


def test_library_exports_every_declared_symbol():
    from sparsesolvers._lib import lib
    l = lib()
    missing = [s for s in _declared_symbols() if not hasattr(l, s)]
    assert not missing, missing


def test_status_strings_are_the_kernelpp_wording():
    """third_party/kernelpp/include/kernelpp/kernel.h:122-133"""
    from sparsesolvers._lib import lib
    l = lib()
    want = {1: "Kernel Failed", 2: "Compute Mode Disabled", 3: "Compute Mode Unavailable",
            4: "Kernel Not Defined", 5: "Cancelled"}
    for code, text in want.items():
        assert l.ssb200_status_string(code).decode() == text


def test_version_matches_reference():
    import sparsesolvers as ss
    assert ss.version() == [0, 8, 8]          # reference CMakeLists.txt:4


def _have_gpu():
    from sparsesolvers._lib import lib
    return lib().ssb200_device_count() > 0


@pytest.mark.skipif(_have_gpu(), reason="checks the no-device behaviour")
def test_no_device_fails_loudly_no_cpu_fallback():
    import sparsesolvers as ss
    with pytest.raises(ss.SolverError, match="Compute Mode Unavailable"):
        ss.Homotopy(np.eye(5, dtype=np.float32))
    with pytest.raises(ss.SolverError, match="Compute Mode Unavailable"):
        ss.Irls(np.eye(5, dtype=np.float64))
    with pytest.raises(ss.SolverError, match="Compute Mode Unavailable"):
        ss.HomotopyGroup(np.eye(5, dtype=np.float32), devices=[0, 1])


def test_product_never_imports_the_oracle():
    """oracle/ is test infrastructure: nothing under sparse-solvers_b200/ may reference it."""
    pkg = os.path.join(ROOT, "sparse-solvers_b200")
    bad = []
    for d, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                txt = open(os.path.join(d, f), errors="ignore").read()
                if re.search(r"^\s*(from|import)\s+oracle\b", txt, flags=re.M) or "ss_oracle" in txt or "libss_ref" in txt:
                    bad.append(os.path.join(d, f))
    assert not bad, bad


def test_argument_checks_mirror_the_binding():
    import sparsesolvers as ss
    with pytest.raises(TypeError):
        ss.Homotopy(np.eye(4, dtype=np.int32))
    with pytest.raises(ValueError):
        ss.Homotopy(np.zeros(4, dtype=np.float32))
