"""ssb200_norm_l1_dev_noalloc_{f32,f64} (include/ssb200.h; no counterpart in the reference):
ss::norm_l1 on a device matrix with its scratch in a static __device__ array, no allocation inside the call.

Checked through the C ABI against the existing entry ssb200_norm_l1_* on the same inputs, two calls in a row (the second on a
smaller matrix, so that what the first call left in the shared scratch array would show), at the tolerance
tests/test_gpu_utilities.py holds ss::norm_l1 to (8 eps, relative).  Collected last on purpose.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

SCRATCH_BYTES = 1 << 20   # SSB200_NOALLOC_SCRATCH_BYTES


def _noalloc(torch, l, dA, stream=None):
    fn = l.ssb200_norm_l1_dev_noalloc_f32 if dA.dtype == torch.float32 else l.ssb200_norm_l1_dev_noalloc_f64
    fn.restype = C.c_int
    s = torch.cuda.current_stream().cuda_stream if stream is None else stream
    return fn(C.c_void_p(dA.data_ptr()), C.c_size_t(dA.shape[0]), C.c_size_t(dA.shape[1]), C.c_size_t(dA.stride(0)),
              C.c_int(0), C.c_void_p(s))


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_noalloc_entry_equals_the_existing_entry_twice_in_a_row(dtype):
    import torch
    import sparsesolvers as ss
    from sparsesolvers._lib import lib
    l = lib()
    rng = np.random.default_rng(21)
    tol = 8 * np.finfo(dtype).eps
    for call, shape in enumerate([(504, 1216), (100, 25)]):        # the second call reuses the scratch the first one filled
        A = rng.normal(0.5, 0.1, size=shape).astype(dtype)
        A[rng.integers(shape[0]), :] *= -1                          # |.|, not a plain sum
        want = A.copy()
        ss.norm_l1(want)                                            # ssb200_norm_l1_*: the existing entry point
        dA = torch.from_numpy(A).cuda()
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated()
        assert _noalloc(torch, l, dA) == 0, l.ssb200_last_error().decode()
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated() == before
        got = dA.cpu().numpy()
        np.testing.assert_allclose(got, want, rtol=tol, atol=0, err_msg=f"call {call} {shape}")
        np.testing.assert_allclose(np.abs(got).sum(axis=0), 1.0, rtol=64 * np.finfo(dtype).eps)


def test_noalloc_entry_on_a_padded_matrix_and_a_side_stream():
    """lda > n (a column block of a wider device matrix: the columns outside stay as they are) and a caller-owned stream"""
    import torch
    import sparsesolvers as ss
    from sparsesolvers._lib import lib
    l = lib()
    rng = np.random.default_rng(22)
    wide = rng.normal(0.5, 0.1, size=(64, 300)).astype(np.float32)
    want = wide.copy()
    ss.norm_l1(want[:, :257])                                       # the existing entry takes the same strided view
    dW = torch.from_numpy(wide).cuda()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    assert _noalloc(torch, l, dW[:, :257], stream=side.cuda_stream) == 0, l.ssb200_last_error().decode()
    side.synchronize()
    np.testing.assert_allclose(dW.cpu().numpy(), want, rtol=8 * np.finfo(np.float32).eps, atol=0)
    # the Python mirror's wrapper of the same entry (torch's current stream)
    dV = torch.from_numpy(wide).cuda()
    ss.norm_l1_noalloc(dV[:, :257])
    torch.cuda.synchronize()
    np.testing.assert_array_equal(dV.cpu().numpy(), dW.cpu().numpy())


def test_noalloc_entry_refuses_what_the_static_scratch_cannot_hold():
    import torch
    from sparsesolvers._lib import lib
    l = lib()
    n = SCRATCH_BYTES // 4 + 1
    dA = torch.ones((1, n), dtype=torch.float32, device="cuda")
    assert _noalloc(torch, l, dA) == 6                              # SSB200_INVALID_ARGUMENT
    msg = l.ssb200_last_error().decode()
    assert "SSB200_NOALLOC_SCRATCH_BYTES" in msg and str(SCRATCH_BYTES) in msg, msg
    torch.cuda.synchronize()
    assert float(dA.sum()) == n                                     # untouched
    dB = torch.ones((2, SCRATCH_BYTES // 4), dtype=torch.float32, device="cuda")   # exactly the limit is served
    assert _noalloc(torch, l, dB) == 0, l.ssb200_last_error().decode()
    torch.cuda.synchronize()
    assert float(dB.min()) == 0.5 and float(dB.max()) == 0.5
