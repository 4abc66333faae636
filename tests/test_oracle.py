"""CPU tests: the plain-C restatement (oracle/ss_oracle.c) pinned against
  (i) committed golden vectors produced by the unmodified reference,
 (ii) the reference itself when oracle/_ref is built (this container),
(iii) the reference's own known-answer tests for this path.
"""
import glob
import os

import numpy as np
import pytest

from conftest import GOLDEN, needs_ref
from oracle import port
import parity


def _load(name):
    return np.load(os.path.join(GOLDEN, name))


HOMOTOPY_GOLDEN = sorted(os.path.basename(p) for p in glob.glob(os.path.join(GOLDEN, "homotopy_*.npz")))
IRLS_GOLDEN = sorted(os.path.basename(p) for p in glob.glob(os.path.join(GOLDEN, "irls_*.npz")) if "_c4" not in p)   # the C4-shape fixture has its own (GPU) test: minutes per signal on the CPU


@pytest.mark.parametrize("name", HOMOTOPY_GOLDEN)
def test_port_homotopy_matches_golden(name):
    g = _load(name)
    A = g["A"]
    if "colmajor" in name:
        A = np.asfortranarray(A)
    X, it, err = port.homotopy(A, g["Y"], float(g["tol"]), int(g["max_iter"]))
    c0 = np.abs(g["Y"].astype(np.float64) @ g["A"].astype(np.float64)).max(axis=1)
    r = parity.compare(X, it, err, g["X"], g["iters"], g["errs"], c0)
    parity.assert_parity(r, g["A"].dtype, name)


@pytest.mark.parametrize("name", IRLS_GOLDEN)
def test_port_irls_matches_golden(name):
    g = _load(name)
    X, it, err, spd = port.irls(g["A"], g["Y"], float(g["tol"]), int(g["max_iter"]))
    assert (it == g["iters"]).all()
    assert (spd == g["spd"]).all()
    np.testing.assert_allclose(err, g["errs"], rtol=1e-4)
    ok = ~np.isnan(g["X"]).any(axis=1)
    assert (np.isnan(X).any(axis=1) == ~ok).all()
    # IRLS weights span ~1/eps, so x is only reproducible to ~1e-3 of its unit sum
    np.testing.assert_allclose(X[ok], g["X"][ok], atol=2e-3)
    assert (np.argmax(X[ok], axis=1) == np.argmax(g["X"][ok], axis=1)).all()


def test_port_worked_example():
    g = _load("worked_5x5.npz")
    X, it, err = port.homotopy(g["A"], g["Y"], 0.05, 5)
    assert it[0] == g["ith"][0]
    np.testing.assert_allclose(X, g["Xh"], rtol=1e-4, atol=1e-6)
    np.testing.assert_allclose(err, g["eh"], rtol=1e-3)
    Xi, iti, ei, si = port.irls(g["A"], g["Y"], 0.05, 5)
    assert iti[0] == g["iti"][0] and bool(si[0]) == bool(g["si"][0])
    np.testing.assert_allclose(Xi, g["Xi"], atol=1e-5)


# ---- the reference's behavioural tests, re-expressed (src/solvers/test_util.h) ----

@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_smoke_identity(dtype):
    """test_util.h:27-55 / bindings/python/tests/test_binding.py:9-20: x == y exactly, iter 1, error 0."""
    N = 5
    A = np.eye(N, dtype=dtype)
    for n in range(N):
        y = np.zeros((1, N), dtype=dtype)
        y[0, n] = 1
        X, it, err = port.homotopy(A, y, 1e-3, N)
        assert (X == y).all() and it[0] == 1 and err[0] == 0.0
        Xi, iti, ei, si = port.irls(A, y, 1e-3, N)
        assert (Xi == y).all() and 1 <= iti[0] <= N


def test_smoke_column_subset():
    """test_util.h:57-92: solver over a strided view (columns 5..9 of a 5x10 buffer)."""
    buf = np.zeros((5, 10), dtype=np.float32)
    buf[:, 5:] = np.eye(5, dtype=np.float32)
    A = buf[:, 5:]
    assert not A.flags["C_CONTIGUOUS"]
    for n in range(5):
        y = np.zeros((1, 5), dtype=np.float32)
        y[0, n] = 1
        X, it, _ = port.homotopy(A, y, 1e-3, 5)
        assert (X == y).all()


def test_noisy_signal():
    """test_util.h:94-126: identity 50, y = e_n + U(0,0.01): exactly one entry above the noise."""
    rng = np.random.default_rng(0)
    N = 50
    A = np.eye(N, dtype=np.float32)
    for n in range(0, N, 7):
        y = rng.uniform(0, 0.01, size=(1, N)).astype(np.float32)
        y[0, n] += 1
        X, it, err = port.homotopy(A, y, 0.01, N)
        assert (X[0] > 0.01).sum() == 1 and np.argmax(X[0]) == n


@pytest.mark.parametrize("shape", [(10, 10), (25, 10), (10, 25)])
def test_permutations(shape):
    """test_util.h:204-257: columns = successive permutations of 1..M + noise; argmax(x) == n."""
    import itertools
    M, N = shape
    rng = np.random.default_rng(0)
    perms = itertools.permutations(range(1, M + 1))
    A = np.stack([np.array(next(perms), dtype=np.float64) for _ in range(N)], axis=1)
    A /= np.abs(A).sum(axis=0, keepdims=True)
    for n in range(N):
        y = (A[:, n] + rng.standard_normal(M) * 1e-4)[None, :]
        X, it, err = port.homotopy(A, y, 1e-3, N)
        assert np.argmax(X[0]) == n


# ---- known-answer tests of the helpers ----

def test_square_permute_known_answers():
    """src/linalg/online_inverse_test.cpp:13-118."""
    A2 = [[1, 2], [3, 4]]
    assert (port.square_permute(A2, 0, 1) == [[4, 3], [2, 1]]).all()
    A3 = [[1, 2, 3], [4, 5, 6], [7, 8, 9]]
    assert (port.square_permute(A3, 1, 2) == [[1, 3, 2], [7, 9, 8], [4, 6, 5]]).all()
    assert (port.square_permute(A3, 0, 2) == [[5, 6, 4], [8, 9, 7], [2, 3, 1]]).all()
    A4 = np.arange(1, 17).reshape(4, 4)
    e13 = [[1, 3, 4, 2], [9, 11, 12, 10], [13, 15, 16, 14], [5, 7, 8, 6]]
    e12 = [[1, 3, 2, 4], [9, 11, 10, 12], [5, 7, 6, 8], [13, 15, 14, 16]]
    assert (port.square_permute(A4, 1, 3) == e13).all()
    assert (port.square_permute(A4, 1, 2) == e12).all()
    assert (port.square_permute(port.square_permute(A4, 1, 3), 3, 1) == A4).all()


def test_online_inverse_identity():
    """online_inverse_test.cpp:186-217: insert then remove 10 identity columns; inverse == I within 1e-4."""
    K = 10
    A = np.eye(K, dtype=np.float32)
    for k in range(1, K + 1):
        inv, idx = port.online_inverse(A, list(range(k)))
        np.testing.assert_allclose(inv, np.eye(k), atol=1e-4)
    ops = list(range(K)) + [-(k + 1) for k in range(K - 1, 2, -1)]
    inv, idx = port.online_inverse(A, ops)
    np.testing.assert_allclose(inv, np.eye(3), atol=1e-4)
    assert idx.tolist() == [0, 1, 2]


@pytest.mark.parametrize("dtype,tol", [(np.float32, 2e-3), (np.float64, 1e-9)])
def test_online_inverse_random(dtype, tol):
    """insert/remove in arbitrary order tracks inv(A_G^T A_G) with G sorted (rank_index.h:64-74)."""
    rng = np.random.default_rng(2)
    A = rng.standard_normal((40, 30)).astype(dtype)
    ops = [7, 3, 21, 12, 0, -(3 + 1), 29, 5, -(21 + 1), 3, 14]
    inv, idx = port.online_inverse(A, ops)
    assert idx.tolist() == sorted(idx.tolist()) == [0, 3, 5, 7, 12, 14, 29]
    G = A[:, idx].astype(np.float64)
    np.testing.assert_allclose(inv, np.linalg.inv(G.T @ G), rtol=tol, atol=tol)


def test_cholesky_known_answers():
    """src/linalg/cholesky_decomposition_test.cpp:16-54."""
    L, ok = port.cholesky(np.array([[0, 1], [1, 0]], dtype=np.float32))
    assert not ok
    A = np.array([[2, 1], [1, 2]], dtype=np.float32)
    L, ok = port.cholesky(A)
    assert ok
    np.testing.assert_allclose(L @ L.T, A, atol=1e-4)
    np.testing.assert_allclose(port.chol_solve(L, [1, -1]), [1, -1], atol=1e-4)


@pytest.mark.parametrize("shape,dtype,tol", [((1, 1), np.float32, 1e-4), ((7, 4), np.float32, 1e-4),
                                              ((50, 50), np.float64, 1e-9), ((100, 20), np.float64, 1e-9)])
def test_qr_properties(shape, dtype, tol):
    """src/linalg/qr_decomposition_test.cpp:36-85: QR == A, Q^T Q == I."""
    rng = np.random.default_rng(1)
    A = rng.random(shape).astype(dtype)
    Q, R = port.qr(A)
    np.testing.assert_allclose(Q @ R, A, atol=tol)
    np.testing.assert_allclose(Q.T @ Q, np.eye(shape[1]), atol=tol)
    assert np.allclose(R, np.triu(R))


# ---- against the reference itself (this container only) ----

@needs_ref
@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_port_vs_reference_homotopy(dtype):
    from oracle import ref
    from sparsesolvers import synth
    A = synth.dictionary(128, 512, dtype=dtype, seed=5)
    Y, _ = synth.sparse_signals(A, 128, 10, seed=6)
    Xr, itr, er = ref.homotopy(A, Y, 1e-3, 40, threads=4)
    X, it, err = port.homotopy(A, Y, 1e-3, 40)
    c0 = np.abs(Y.astype(np.float64) @ A.astype(np.float64)).max(axis=1)
    r = parity.compare(X, it, err, Xr, itr, er, c0)
    parity.assert_parity(r, dtype, "port vs reference")


@needs_ref
def test_port_vs_reference_irls():
    from oracle import ref
    rng = np.random.default_rng(9)
    A = (rng.standard_normal((96, 24)) / np.sqrt(96))
    X0 = np.zeros((6, 24))
    for b in range(6):
        X0[b, rng.choice(24, 2, replace=False)] = 1 + np.abs(rng.standard_normal(2))
    Y = X0 @ A.T
    Xr, itr, er, sr = ref.irls(A, Y, 1e-3, 50)
    X, it, e, s = port.irls(A, Y, 1e-3, 50)
    assert (it == itr).all() and (s == sr).all()
    np.testing.assert_allclose(e, er, rtol=1e-6)


@needs_ref
def test_reference_norm_l1_and_reconstruct_known_answers():
    """src/linalg/norms_test.cpp:9-24 through the reference's public ss::norm_l1, and y = A x through ss::reconstruct_signal
    (include/ss/ss.h:79-93): pins the oracle/_ref entry points the GPU utilities are tested against."""
    from oracle import ref
    A = np.array([[1, 2, 0], [3, 4, 1]], dtype=np.float32)
    np.testing.assert_allclose(ref.norm_l1(A), [[0.25, 0.3333, 0], [0.75, 0.6667, 1]], rtol=0, atol=1e-4)
    np.testing.assert_allclose(ref.reconstruct_signal(A, np.array([1, 1, 1], dtype=np.float32)), [3, 8])
    Ad = np.arange(12, dtype=np.float64).reshape(3, 4)
    np.testing.assert_allclose(ref.reconstruct_signal(Ad, np.array([1.0, 0, -1, 2])), Ad @ np.array([1.0, 0, -1, 2]))
