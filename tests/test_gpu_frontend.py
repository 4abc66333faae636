"""The batch front-end for device-resident data: torch CUDA tensors (or any DLPack /
__cuda_array_interface__ producer) go through the device-pointer entries of the C ABI without touching the host, and give
exactly what the host-array path gives -- which the other test modules hold against the oracle."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def case():
    import sparsesolvers as ss
    from sparsesolvers import synth
    A = synth.dictionary(128, 512, dtype=np.float32, seed=31)
    Y, _ = synth.sparse_signals(A, 200, 9, seed=32)
    s = ss.Homotopy(A)
    X, it, err = s.solve_batch(Y, 1e-3, 40)
    return ss, A, Y, s, X, it, err


def test_cuda_signals_equal_host_path(case):
    import torch
    ss, A, Y, s, X, it, err = case
    dY = torch.from_numpy(Y).cuda()
    dX, dit, derr = s.solve_batch(dY, 1e-3, 40)
    assert dX.is_cuda and dX.dtype == torch.float32 and dit.dtype == torch.int32
    assert (dit.cpu().numpy().astype(np.uint32) == it).all()
    assert np.array_equal(dX.cpu().numpy(), X)              # a signal's arithmetic does not depend on its batch
    assert np.array_equal(derr.cpu().numpy(), err)


def test_oracle_parity_through_the_device_front_end(case):
    import torch
    import parity
    from gpu_util import c0_inf, cpu_reference
    ss, A, Y, s, X, it, err = case
    (Xr, itr, er), _ = cpu_reference(A, Y, 1e-3, 40)
    dX, dit, derr = s.solve_batch(torch.from_numpy(Y).cuda(), 1e-3, 40)
    r = parity.compare(dX.cpu().numpy(), dit.cpu().numpy().astype(np.uint32), derr.cpu().numpy(), Xr, itr, er, c0_inf(A, Y))
    parity.assert_parity(r, np.float32, "device front-end")


def test_strided_signals_and_out(case):
    import torch
    ss, A, Y, s, X, it, err = case
    wide = torch.zeros((Y.shape[0], 160), dtype=torch.float32, device="cuda")
    wide[:, :128] = torch.from_numpy(Y).cuda()
    out = torch.full((Y.shape[0], 512), 7.0, dtype=torch.float32, device="cuda")
    dX, dit, _ = s.solve_batch(wide[:, :128], 1e-3, 40, out=out)      # row stride 160, no copy
    assert dX.data_ptr() == out.data_ptr()
    assert np.array_equal(out.cpu().numpy(), X)
    dXt, _, _ = s.solve_batch(torch.from_numpy(np.ascontiguousarray(Y.T)).cuda().T, 1e-3, 40)   # inner stride != 1: repacked
    assert np.array_equal(dXt.cpu().numpy(), X)


def test_sparse_result_form(case):
    import torch
    ss, A, Y, s, X, it, err = case
    (idx, val, cnt), dit, _ = s.solve_batch(torch.from_numpy(Y).cuda(), 1e-3, 40, sparse=True)
    idx, val, cnt = idx.cpu().numpy(), val.cpu().numpy(), cnt.cpu().numpy()
    assert idx.shape == (200, 41) and (cnt == (X != 0).sum(axis=1)).all() and cnt.max() <= 41
    dense = np.zeros_like(X)
    for b in range(X.shape[0]):
        k = cnt[b]
        assert (np.diff(idx[b, :k]) > 0).all() and (idx[b, k:] == -1).all() and (val[b, k:] == 0).all()
        dense[b, idx[b, :k]] = val[b, :k]
    assert np.array_equal(dense, X)


def test_device_resident_dictionary_and_reconstruct(case):
    import torch
    ss, A, Y, s, X, it, err = case
    sd = ss.Homotopy(torch.from_numpy(A).cuda())
    assert sd.shape == (128, 512) and sd.dtype == np.float32
    dX, dit, _ = sd.solve_batch(torch.from_numpy(Y).cuda(), 1e-3, 40)
    assert np.array_equal(dX.cpu().numpy(), X)
    sf = ss.Homotopy(torch.from_numpy(np.asfortranarray(A)).cuda())   # column-major device tensor: strides pass through
    assert np.array_equal(sf.solve_batch(torch.from_numpy(Y).cuda(), 1e-3, 40)[0].cpu().numpy(), X)
    R = sd.reconstruct(dX)                                            # ss::reconstruct_signal, batched on the device
    assert R.is_cuda and np.array_equal(R.cpu().numpy(), s.reconstruct(X))
    good = err <= 1e-3
    assert np.abs(R.cpu().numpy() - Y)[good].max() < 5e-3


def test_float64_and_argument_checks(case):
    import torch
    ss, A, Y, s, X, it, err = case
    A64, Y64 = A.astype(np.float64), Y.astype(np.float64)
    s64 = ss.Homotopy(A64)
    Xh, ith, _ = s64.solve_batch(Y64[:50], 1e-3, 40)
    Xd, itd, _ = s64.solve_batch(torch.from_numpy(Y64[:50]).cuda(), 1e-3, 40)
    assert np.array_equal(Xd.cpu().numpy(), Xh) and (itd.cpu().numpy() == ith).all()
    with pytest.raises(TypeError):
        s.solve_batch(torch.from_numpy(Y64).cuda(), 1e-3, 40)         # noconvert: dtype must match (binding.cpp:108)
    with pytest.raises(ValueError):
        s.solve_batch(torch.zeros((4, 100), dtype=torch.float32, device="cuda"), 1e-3, 40)
    with pytest.raises(ValueError):
        s.solve_batch(Y, 1e-3, 40, sparse=True, out=np.zeros((200, 512), dtype=np.float32))   # out= belongs to the dense form
    with pytest.raises(TypeError):
        ss.Irls(torch.from_numpy(A.T.copy()).cuda())
    X0, it0, e0 = s.solve_batch(torch.zeros((0, 128), dtype=torch.float32, device="cuda"), 1e-3, 40)
    assert X0.shape == (0, 512) and it0.numel() == 0


@pytest.mark.parametrize("pad", [0, 24])
def test_host_entry_sparse_return_equals_device_entry(pad):
    """Large host batches return their coefficients compacted over PCIe and scattered into the caller's zero-filled X
    (solve_batch_host: sparse_out); the caller must see exactly the dense X of the device-pointer entry, in a padded X too,
    and whatever the buffer held before."""
    import torch
    import sparsesolvers as ss
    from sparsesolvers import synth
    A = synth.dictionary(128, 1024, dtype=np.float32, seed=41)
    Y, _ = synth.sparse_signals(A, 8192 + 77, 8, seed=42)
    s = ss.Homotopy(A)
    dX, dit, derr = s.solve_batch(torch.from_numpy(Y).cuda(), 1e-3, 40)
    wide = np.full((Y.shape[0], 1024 + pad), np.float32(3.5))
    X, it, err = s.solve_batch(Y, 1e-3, 40, out=wide[:, :1024])
    assert X.base is wide or X is wide or np.shares_memory(X, wide)
    assert np.array_equal(X, dX.cpu().numpy())
    assert (it == dit.cpu().numpy().astype(np.uint32)).all() and np.array_equal(err, derr.cpu().numpy())
    if pad:
        assert (wide[:, 1024:] == np.float32(3.5)).all()       # columns outside the view are not touched


@pytest.mark.parametrize("batch", [3, 200, 8192 + 77])
def test_host_entry_sparse_result(batch):
    """ssb200_homotopy_solve_batch_sparse_*: the host-pointer entry hands the coefficients back as (index, value) x cap + count per
    signal -- no dense row is formed on the host; scattered back they are the dense result, entry for entry."""
    import sparsesolvers as ss
    from sparsesolvers import synth
    A = synth.dictionary(128, 1024, dtype=np.float32, seed=41)
    Y, _ = synth.sparse_signals(A, batch, 8, seed=42)
    s = ss.Homotopy(A)
    X, it, err = s.solve_batch(Y, 1e-3, 40)
    (idx, val, cnt), its, errs = s.solve_batch(Y, 1e-3, 40, sparse=True)
    assert idx.shape == (batch, 41) and (its == it).all() and np.array_equal(errs, err)
    assert (cnt == (X != 0).sum(axis=1)).all()
    dense = np.zeros_like(X)
    for b in range(batch):
        k = cnt[b]
        assert (np.diff(idx[b, :k]) > 0).all()
        dense[b, idx[b, :k]] = val[b, :k]
    assert np.array_equal(dense, X)
