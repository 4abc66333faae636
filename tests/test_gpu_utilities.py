"""ss::reconstruct_signal and ss::norm_l1 on the device (include/ss/ss.h:79-93, src/lib.cpp:78-112,
src/linalg/norms.h:22-33) against the UNMODIFIED reference's own entry points (oracle/_ref: ssref_reconstruct_*,
ssref_norm_l1_*) and the reference's known answers (src/linalg/norms_test.cpp:9-37)."""
import numpy as np
import pytest

from conftest import needs_ref

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ss():
    import sparsesolvers
    return sparsesolvers


def test_norm_l1_known_answers(ss):
    """norms_test.cpp:9-24 (matrix) and :26-37 (the same vector as a one-column matrix), 4 decimals as the reference checks them"""
    A = np.array([[1, 2, 0], [3, 4, 1]], dtype=np.float32)
    ss.norm_l1(A)
    np.testing.assert_allclose(A, [[0.25, 0.3333, 0], [0.75, 0.6667, 1]], rtol=0, atol=1e-4)
    x = np.array([[1], [2], [3], [4], [5], [0]], dtype=np.float32)
    ss.norm_l1(x)
    np.testing.assert_allclose(x[:, 0], [0.06667, 0.1333, 0.2, 0.2666, 0.3333, 0], rtol=0, atol=1e-4)


@needs_ref
@pytest.mark.parametrize("dtype,shape", [(np.float32, (100, 25)), (np.float32, (25, 100)), (np.float64, (504, 1216)), (np.float32, (7, 3))])
def test_norm_l1_equals_the_reference(ss, dtype, shape):
    from oracle import ref
    rng = np.random.default_rng(3)
    A = rng.normal(0.5, 0.1, size=shape).astype(dtype)
    want = ref.norm_l1(A)
    got = A.copy()
    ss.norm_l1(got)
    np.testing.assert_allclose(got, want, rtol=8 * np.finfo(dtype).eps, atol=0)
    view = np.asfortranarray(A)                                   # a column-major view goes through the same strides contract
    ss.norm_l1(view)
    np.testing.assert_allclose(view, want, rtol=8 * np.finfo(dtype).eps, atol=0)


@needs_ref
@pytest.mark.parametrize("dtype,shape,k", [(np.float32, (100, 25), 5), (np.float32, (256, 1024), 16), (np.float64, (128, 512), 10), (np.float32, (48, 100), 100)])
def test_reconstruct_signal_equals_the_reference(ss, dtype, shape, k):
    from oracle import ref
    rng = np.random.default_rng(4)
    A = (rng.standard_normal(shape) / np.sqrt(shape[0])).astype(dtype)
    X = np.zeros((6, shape[1]), dtype=dtype)
    for b in range(6):
        X[b, rng.choice(shape[1], k, replace=False)] = rng.standard_normal(k)
    s = ss.Homotopy(A)
    Y = s.reconstruct(X)
    scale = np.abs(A).max() * np.abs(X).sum(axis=1)                # summation order differs (OpenBLAS gemv vs one thread per row)
    for b in range(6):
        want = ref.reconstruct_signal(A, X[b])
        np.testing.assert_allclose(Y[b], want, rtol=0, atol=8 * np.finfo(dtype).eps * scale[b])
        np.testing.assert_allclose(ss.reconstruct_signal(A, X[b]), Y[b], rtol=0, atol=0)


def test_reconstruct_on_device_tensors(ss):
    import torch
    rng = np.random.default_rng(5)
    A = (rng.standard_normal((64, 256)) / 8).astype(np.float32)
    X = np.zeros((9, 256), dtype=np.float32)
    X[:, ::17] = rng.standard_normal((9, 16)).astype(np.float32)
    s = ss.Homotopy(A)
    Yh = s.reconstruct(X)
    Yd = s.reconstruct(torch.from_numpy(X).cuda())
    np.testing.assert_array_equal(Yd.cpu().numpy(), Yh)
