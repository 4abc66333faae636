"""ctypes loader for the plain-C restatement (oracle/ss_oracle.c).

TEST INFRASTRUCTURE ONLY — see the header of ss_oracle.c.  Imported by tests/,
__graft_entry__.smoke() and bench.py's CPU legs; never by the product.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libss_oracle.so")
_lib = None


def build(force=False):
    src = [os.path.join(_HERE, "ss_oracle.c"), os.path.join(_HERE, "ss_oracle_impl.h")]
    if not force and os.path.exists(_SO) and all(os.path.getmtime(_SO) >= os.path.getmtime(s) for s in src):
        return _SO
    os.makedirs(os.path.dirname(_SO), exist_ok=True)
    subprocess.check_call(["gcc", "-O2", "-fopenmp", "-fPIC", "-shared", "-o", _SO, src[0], "-lm"])
    return _SO


def _load():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_SO)
    return _lib


def _sfx(dtype):
    return {np.dtype(np.float32): ("f32", C.c_float), np.dtype(np.float64): ("f64", C.c_double)}[np.dtype(dtype)]


def _p(a):
    return C.c_void_p(a.ctypes.data)


def homotopy(A, Y, tol, max_iter):
    lib = _load()
    A = np.asarray(A)
    Y = np.ascontiguousarray(np.atleast_2d(Y), dtype=A.dtype)
    sfx, ct = _sfx(A.dtype)
    m, n = A.shape
    B = Y.shape[0]
    X = np.zeros((B, n), dtype=A.dtype)
    iters = np.zeros(B, dtype=np.uint32)
    errs = np.zeros(B, dtype=np.float64)
    rs, cs = A.strides[0] // A.itemsize, A.strides[1] // A.itemsize
    getattr(lib, "ssor_homotopy_batch_" + sfx)(
        _p(A), C.c_size_t(m), C.c_size_t(n), C.c_ssize_t(rs), C.c_ssize_t(cs),
        _p(Y), C.c_size_t(B), C.c_size_t(m), ct(tol), C.c_uint32(max_iter),
        _p(X), C.c_size_t(n), _p(iters), _p(errs))
    return X, iters, errs


def irls(A, Y, tol, max_iter):
    lib = _load()
    A = np.asarray(A)
    Y = np.ascontiguousarray(np.atleast_2d(Y), dtype=A.dtype)
    sfx, ct = _sfx(A.dtype)
    m, n = A.shape
    B = Y.shape[0]
    X = np.zeros((B, n), dtype=A.dtype)
    iters = np.zeros(B, dtype=np.uint32)
    errs = np.zeros(B, dtype=np.float64)
    spd = np.zeros(B, dtype=np.uint8)
    rs, cs = A.strides[0] // A.itemsize, A.strides[1] // A.itemsize
    getattr(lib, "ssor_irls_" + sfx)(
        _p(A), C.c_size_t(m), C.c_size_t(n), C.c_ssize_t(rs), C.c_ssize_t(cs),
        _p(Y), C.c_size_t(B), C.c_size_t(m), ct(tol), C.c_uint32(max_iter),
        _p(X), C.c_size_t(n), _p(iters), _p(errs), _p(spd))
    return X, iters, errs, spd.astype(bool)


def qr(A):
    lib = _load()
    A = np.ascontiguousarray(A)
    sfx, _ = _sfx(A.dtype)
    m, n = A.shape
    Q = np.zeros((m, n), dtype=A.dtype)
    R = np.zeros((n, n), dtype=A.dtype)
    getattr(lib, "ssor_qr_" + sfx)(_p(A), C.c_size_t(m), C.c_size_t(n), C.c_ssize_t(n), C.c_ssize_t(1), _p(Q), _p(R))
    return Q, R


def cholesky(G):
    lib = _load()
    G = np.ascontiguousarray(G)
    sfx, _ = _sfx(G.dtype)
    n = G.shape[0]
    L = np.zeros((n, n), dtype=G.dtype)
    fn = getattr(lib, "ssor_cholesky_" + sfx)
    fn.restype = C.c_int
    ok = fn(_p(G), C.c_size_t(n), _p(L))
    return L, bool(ok)


def chol_solve(L, b):
    lib = _load()
    L = np.ascontiguousarray(L)
    sfx, _ = _sfx(L.dtype)
    x = np.array(b, dtype=L.dtype, copy=True)
    getattr(lib, "ssor_chol_solve_" + sfx)(_p(L), C.c_size_t(L.shape[0]), _p(x))
    return x


def online_inverse(A, ops):
    """ops: sequence of ints; c >= 0 toggles column c in (insert), -(c+1) toggles it out."""
    lib = _load()
    A = np.ascontiguousarray(A)
    sfx, _ = _sfx(A.dtype)
    m, n = A.shape
    ops = np.asarray(ops, dtype=np.int32)
    inv = np.zeros(len(ops) * len(ops) + 1, dtype=A.dtype)
    idx = np.zeros(len(ops) + 1, dtype=np.uint32)
    fn = getattr(lib, "ssor_online_inverse_" + sfx)
    fn.restype = C.c_size_t
    k = fn(_p(A), C.c_size_t(m), C.c_size_t(n), _p(ops), C.c_size_t(len(ops)), _p(inv), _p(idx))
    return inv[:k * k].reshape(k, k).copy(), idx[:k].copy()


def square_permute(A, src, dest):
    lib = _load()
    A = np.array(A, dtype=np.float32, copy=True)
    lib.ssor_square_permute_f32(_p(A), C.c_size_t(A.shape[0]), C.c_size_t(src), C.c_size_t(dest))
    return A
