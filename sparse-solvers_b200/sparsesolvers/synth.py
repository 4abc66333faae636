"""Synthetic workloads for the homotopy / IRLS hot path.

The reference prescribes no inputs (its tests draw from xtensor's RNG in-test,
src/solvers/test_util.h:136-197), so these generators are this repo's
definitions; the SAME host arrays are fed to the oracle and to the GPU path.

  A   : i.i.d. N(0, 1/M), row-major (M, N)           -- unit expected column norm
  x0  : k distinct uniform positions, values s*(1+|g|), g~N(0,1), s=+-1
  y   : A @ x0 accumulated in float64, rounded to dtype
  sign normalisation: negate (x0, y) when the first selected correlation
        c[argmax|c|] (c = A^T y) is negative.  The reference initialises its
        direction with sign(|c|) (src/solvers/homotopy-cpu.cpp:223-227), so an
        un-normalised signal takes the divergent path to max_iter; the "quirk"
        sets keep those signals to check that behaviour is reproduced.
"""
import numpy as np

CONFIGS = {
    # name: (M, N, k, dtype, batch, tol, max_iter)          BASELINE.json configs[]
    "C1": dict(m=256, n=1024, k=16, dtype="float32", batch=1024, tol=1e-3, max_iter=48),
    "C2": dict(m=512, n=2048, k=32, dtype="float32", batch=65536, tol=1e-3, max_iter=80),
    "C3": dict(m=504, n=1216, k=32, dtype="float32", batch=262144, tol=1e-3, max_iter=100),
    "C5": dict(m=4096, n=16384, k=128, dtype="float32", batch=8192, tol=1e-3, max_iter=272),
}


def dictionary(m, n, dtype=np.float32, seed=1234):
    rng = np.random.default_rng(seed)
    return (rng.standard_normal((m, n)) / np.sqrt(m)).astype(dtype)


def sparse_signals(A, batch, k, seed=4321, normalise_sign=True, chunk=4096):
    """Returns (Y (batch, m), X0 (batch, n)) in A.dtype."""
    m, n = A.shape
    rng = np.random.default_rng(seed)
    A64 = A.astype(np.float64)
    Y = np.empty((batch, m), dtype=A.dtype)
    X0 = np.zeros((batch, n), dtype=A.dtype)
    for lo in range(0, batch, chunk):
        hi = min(batch, lo + chunk)
        b = hi - lo
        # k distinct positions per signal: k smallest of n uniform keys
        keys = rng.random((b, n), dtype=np.float32)
        idx = np.argpartition(keys, k, axis=1)[:, :k]
        g = rng.standard_normal((b, k))
        s = np.where(rng.random((b, k)) < 0.5, -1.0, 1.0)
        x0 = np.zeros((b, n), dtype=np.float64)
        np.put_along_axis(x0, idx, s * (1.0 + np.abs(g)), axis=1)
        x0 = x0.astype(A.dtype).astype(np.float64)
        y = x0 @ A64.T
        if normalise_sign:
            c = y.astype(A.dtype).astype(np.float64) @ A64
            j = np.argmax(np.abs(c), axis=1)
            flip = c[np.arange(b), j] < 0
            y[flip] *= -1.0
            x0[flip] *= -1.0
        Y[lo:hi] = y.astype(A.dtype)
        X0[lo:hi] = x0.astype(A.dtype)
    return Y, X0


def face_recognition(m=504, n=1216, batch=4096, classes=38, corruption=0.10, seed=77,
                     dtype=np.float32, chunk=4096):
    """C3 shape (Yang et al.): unit-l2 columns, x0 on one class block with non-negative
    weights summing to 1, 10 % of the entries of y replaced by U(-1,1)*|y|_inf."""
    rng = np.random.default_rng(seed)
    A = rng.standard_normal((m, n)) / np.sqrt(m)
    A /= np.linalg.norm(A, axis=0, keepdims=True)
    A = A.astype(dtype)
    A64 = A.astype(np.float64)
    per = n // classes
    Y = np.empty((batch, m), dtype=dtype)
    labels = np.empty(batch, dtype=np.int64)
    ncorrupt = int(np.floor(corruption * m))
    for lo in range(0, batch, chunk):
        hi = min(batch, lo + chunk)
        b = hi - lo
        cls = rng.integers(0, classes, size=b)
        w = rng.random((b, per))
        w /= w.sum(axis=1, keepdims=True)
        x0 = np.zeros((b, n))
        cols = cls[:, None] * per + np.arange(per)[None, :]
        np.put_along_axis(x0, cols, w, axis=1)
        y = x0 @ A64.T
        if ncorrupt:
            pos = np.argpartition(rng.random((b, m), dtype=np.float32), ncorrupt, axis=1)[:, :ncorrupt]
            amp = np.abs(y).max(axis=1, keepdims=True)
            np.put_along_axis(y, pos, rng.uniform(-1, 1, size=(b, ncorrupt)) * amp, axis=1)
        c = y.astype(dtype).astype(np.float64) @ A64
        j = np.argmax(np.abs(c), axis=1)
        flip = c[np.arange(b), j] < 0
        y[flip] *= -1.0
        Y[lo:hi] = y.astype(dtype)
        labels[lo:hi] = cls
    return A, Y, labels


def thresholded_support(X, rel):
    """Support definition used for parity: |x_j| > rel * max|x|."""
    X = np.atleast_2d(X)
    mx = np.abs(X).max(axis=1, keepdims=True)
    return np.abs(X) > rel * mx
