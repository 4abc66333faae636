"""ctypes loader for oracle/_ref/libss_ref.so — the UNMODIFIED reference solver
(built by oracle/ref_build/Makefile from /root/reference/src, never copied).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py.  The product
(sparse-solvers_b200/) must never import this module.

Entry points mirror ref_capi.cpp; they call ss::homotopy<T>::solve /
ss::irls<T>::solve (reference include/ss/ss.h:107-115) once per signal.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_REF_DIR = os.path.join(_HERE, "_ref")
_lib = None
_fwd = None


def available():
    return os.path.exists(os.path.join(_REF_DIR, "libss_ref.so")) and \
        os.path.exists(os.path.join(_REF_DIR, "libcblas_fwd.so"))


def _load():
    global _lib, _fwd
    if _lib is not None:
        return _lib
    if not available():
        raise RuntimeError("oracle/_ref is not built: run `make -C oracle/ref_build` "
                           "(needs /root/reference) or __graft_entry__.build()")
    # the reference dlopen()s /root/repo/oracle/_ref/libcblas_fwd.so itself; when the
    # repo lives elsewhere, preload the shim so dlopen() by soname/path still resolves.
    _fwd = C.CDLL(os.path.join(_REF_DIR, "libcblas_fwd.so"), mode=C.RTLD_GLOBAL)
    want = "/root/repo/oracle/_ref/libcblas_fwd.so"
    if not os.path.exists(want):
        raise RuntimeError(f"{want} missing: the reference late-binds BLAS from that path")
    _lib = C.CDLL(os.path.join(_REF_DIR, "libss_ref.so"))
    _fwd.ssref_blas_config.restype = C.c_char_p
    _fwd.ssref_blas_get_num_threads.restype = C.c_int
    return _lib


def blas_config():
    _load()
    return _fwd.ssref_blas_config().decode()


def set_blas_threads(n):
    _load()
    _fwd.ssref_blas_set_num_threads(int(n))


def get_blas_threads():
    _load()
    return int(_fwd.ssref_blas_get_num_threads())


def _ct(dtype):
    return {np.dtype(np.float32): (C.c_float, "f32"), np.dtype(np.float64): (C.c_double, "f64")}[np.dtype(dtype)]


def _strides_elems(A):
    return A.strides[0] // A.itemsize, A.strides[1] // A.itemsize


def homotopy(A, Y, tol, max_iter, threads=1):
    """Reference ss::homotopy on each row of Y.  A: (m, n) any strides; Y: (B, m) C-contiguous.
    Returns X (B, n), iters (B,) uint32, errs (B,) float64."""
    lib = _load()
    A = np.asarray(A)
    Y = np.ascontiguousarray(np.atleast_2d(Y), dtype=A.dtype)
    ct, sfx = _ct(A.dtype)
    m, n = A.shape
    B = Y.shape[0]
    assert Y.shape[1] == m
    X = np.zeros((B, n), dtype=A.dtype)
    iters = np.zeros(B, dtype=np.uint32)
    errs = np.zeros(B, dtype=np.float64)
    rs, cs = _strides_elems(A)
    fn = getattr(lib, "ssref_homotopy_" + sfx)
    fn.restype = C.c_int
    rc = fn(C.c_void_p(A.ctypes.data), C.c_size_t(m), C.c_size_t(n), C.c_ssize_t(rs), C.c_ssize_t(cs),
            C.c_void_p(Y.ctypes.data), C.c_size_t(B), C.c_size_t(m), ct(tol), C.c_uint32(max_iter),
            C.c_void_p(X.ctypes.data), C.c_size_t(n), C.c_void_p(iters.ctypes.data),
            C.c_void_p(errs.ctypes.data), C.c_int(threads))
    if rc != 0:
        raise RuntimeError("reference homotopy returned an error")
    return X, iters, errs


def irls(A, Y, tol, max_iter, threads=1):
    """Reference ss::irls on each row of Y.  Returns X, iters, errs (=eps), spd_failure (B,) bool."""
    lib = _load()
    A = np.asarray(A)
    Y = np.ascontiguousarray(np.atleast_2d(Y), dtype=A.dtype)
    ct, sfx = _ct(A.dtype)
    m, n = A.shape
    B = Y.shape[0]
    X = np.zeros((B, n), dtype=A.dtype)
    iters = np.zeros(B, dtype=np.uint32)
    errs = np.zeros(B, dtype=np.float64)
    spd = np.zeros(B, dtype=np.uint8)
    rs, cs = _strides_elems(A)
    fn = getattr(lib, "ssref_irls_" + sfx)
    fn.restype = C.c_int
    rc = fn(C.c_void_p(A.ctypes.data), C.c_size_t(m), C.c_size_t(n), C.c_ssize_t(rs), C.c_ssize_t(cs),
            C.c_void_p(Y.ctypes.data), C.c_size_t(B), C.c_size_t(m), ct(tol), C.c_uint32(max_iter),
            C.c_void_p(X.ctypes.data), C.c_size_t(n), C.c_void_p(iters.ctypes.data),
            C.c_void_p(errs.ctypes.data), C.c_void_p(spd.ctypes.data), C.c_int(threads))
    if rc != 0:
        raise RuntimeError("reference irls returned an error")
    return X, iters, errs, spd.astype(bool)


def reconstruct_signal(A, x):
    """Reference ss::reconstruct_signal (include/ss/ss.h:79-86, src/lib.cpp:78-104): y = A x.  A (m, n) row-major."""
    lib = _load()
    A = np.ascontiguousarray(A)
    ct, sfx = _ct(A.dtype)
    m, n = A.shape
    x = np.ascontiguousarray(x, dtype=A.dtype)
    y = np.zeros(m, dtype=A.dtype)
    fn = getattr(lib, "ssref_reconstruct_" + sfx)
    fn.restype = None
    fn(C.c_void_p(A.ctypes.data), C.c_size_t(m), C.c_size_t(n), C.c_void_p(x.ctypes.data), C.c_void_p(y.ctypes.data))
    return y


def norm_l1(A):
    """Reference ss::norm_l1 (include/ss/ss.h:88-93, src/linalg/norms.h:22-33) on a copy of A; returns the normalised copy."""
    lib = _load()
    A = np.array(A, order="C", copy=True)
    ct, sfx = _ct(A.dtype)
    fn = getattr(lib, "ssref_norm_l1_" + sfx)
    fn.restype = None
    fn(C.c_void_p(A.ctypes.data), C.c_size_t(A.shape[0]), C.c_size_t(A.shape[1]))
    return A
