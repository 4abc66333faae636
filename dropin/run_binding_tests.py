"""The cases of the reference's bindings/python/tests/test_binding.py (:9-77), run against the reference's OWN pybind11
module built on top of libssb200 (dropin/Makefile).  /root/reference does not exist on the GPU box, so the cases are
restated here; when it is present the reference's file itself is run too.  Exit code 0 = all passed."""
import os
import sys
import unittest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "_build"))
import numpy as np           # noqa: E402
import sparsesolvers as ss   # noqa: E402  (the pybind11 module, not the ctypes mirror)

assert ss.__file__.startswith(os.path.join(HERE, "_build")), ss.__file__


def _test_smoke(S, N, T):
    A = np.identity(N, dtype=T)
    solver = S(A)
    for n in range(0, N - 1):
        signal = np.zeros(N, dtype=T)
        signal[n] = 1
        x, info = solver.solve(signal)
        assert np.array_equal(signal, x)
        assert info.solution_error == 0
        assert info.iter == 1


class HomotopySolverTest(unittest.TestCase):
    def test_smoke_f32(self):
        _test_smoke(ss.Homotopy, 5, np.float32)

    def test_smoke_f64(self):
        _test_smoke(ss.Homotopy, 5, np.float64)

    def test_row_subset(self):
        A = np.random.rand(10, 5) * 0.1
        A_sub = A[:5, :]
        A_sub[:, 0] = 1
        x, info = ss.Homotopy(A_sub).solve(np.ones(5))
        assert len(x) == 5 and np.linalg.norm(x, 0) == 1

    def test_col_subset(self):
        A = np.random.rand(10, 5) * 0.1
        A[:, 0] = 1
        A[:, 3] = 1
        x, info = ss.Homotopy(A[:, 2:]).solve(np.ones(10))
        assert len(x) == 3 and np.argmax(x) == 1

    def test_transpose(self):
        A = np.random.rand(5, 10) * 0.1
        A[3, :] = 1
        x, info = ss.Homotopy(A.T).solve(np.ones(10))
        assert len(x) == 5 and np.argmax(x) == 3

    def test_view_edited_in_place(self):
        """the reference state is a non-owning view (policies.h:42): edits to A between solves must be seen"""
        A = np.identity(4)
        s = ss.Homotopy(A)
        y = np.zeros(4)
        y[1] = 1
        assert s.solve(y)[0][1] == 1
        A[1, 1] = 2
        assert s.solve(y)[0][1] == 0.5


class IrlsSolverTest(unittest.TestCase):
    def test_smoke_f32(self):
        _test_smoke(ss.Irls, 5, np.float32)

    def test_smoke_f64(self):
        _test_smoke(ss.Irls, 5, np.float64)


if __name__ == "__main__":
    print("[sparsesolvers] version={}".format(ss.version()))
    suite = unittest.defaultTestLoader.loadTestsFromModule(sys.modules[__name__])
    ref_tests = "/root/reference/bindings/python/tests"
    if os.path.isdir(ref_tests):
        sys.path.insert(0, ref_tests)
        suite.addTests(unittest.defaultTestLoader.loadTestsFromName("test_binding"))
    ok = unittest.TextTestRunner(verbosity=1).run(suite).wasSuccessful()
    sys.exit(0 if ok else 1)
