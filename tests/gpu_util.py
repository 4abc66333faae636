"""Helpers shared by the -m gpu parity tests."""
import numpy as np

import parity


def c0_inf(A, Y):
    return np.abs(np.asarray(Y, dtype=np.float64) @ np.asarray(A, dtype=np.float64)).max(axis=1)


def check_homotopy(solver, A, Y, tol, max_iter, Xr, itr, er, what, noise_rate=0.02):
    X, it, err = solver.solve_batch(Y, tol, max_iter)
    r = parity.compare(X, it, err, Xr, itr, er, c0_inf(A, Y))
    return parity.assert_parity(r, A.dtype, what, noise_rate=noise_rate), (X, it, err)


def _next_perm(v):
    v = list(v)
    i = len(v) - 2
    while i >= 0 and v[i] >= v[i + 1]:
        i -= 1
    if i < 0:
        return v[::-1]
    j = len(v) - 1
    while v[j] <= v[i]:
        j -= 1
    v[i], v[j] = v[j], v[i]
    v[i + 1:] = reversed(v[i + 1:])
    return v


def _permute(v, n):
    for _ in range(n):
        v = _next_perm(v)
    return v


def permutation_case(M, N, dtype, noise, skip, seed=0):
    """The reference's permutations_test inputs (src/solvers/test_util.h:204-257): column n of A is the
    (n*skip)-th next_permutation of 1..M plus N(0, noise) sensing noise; signal n is the same permutation
    plus N(0, noise) signal noise.  The reference draws from xtensor's RNG; numpy's is used here, so the
    expected answers are taken from the CPU oracle on the same draws, not assumed."""
    rng = np.random.default_rng(seed)
    col = _permute(list(range(1, M + 1)), skip)
    A = rng.normal(0, noise, size=(M, N))
    c = list(col)
    for n in range(N):
        A[:, n] += c
        c = _permute(c, skip)
    sigs, c = [], list(col)
    for n in range(N):
        sigs.append(np.array(c, dtype=np.float64) + rng.normal(0, noise, size=M))
        c = _permute(c, skip)
    return A.astype(dtype), np.array(sigs).astype(dtype)
