"""Helpers shared by the -m gpu parity tests."""
import numpy as np

import parity


def c0_inf(A, Y):
    return np.abs(np.asarray(Y, dtype=np.float64) @ np.asarray(A, dtype=np.float64)).max(axis=1)


def check_homotopy(solver, A, Y, tol, max_iter, Xr, itr, er, what, noise_rate=0.02):
    X, it, err = solver.solve_batch(Y, tol, max_iter)
    r = parity.compare(X, it, err, Xr, itr, er, c0_inf(A, Y))
    return parity.assert_parity(r, A.dtype, what, noise_rate=noise_rate), (X, it, err)
