"""The multi-GPU group entry of the C ABI (ssb200_group_*: SURVEY.md 8(b) last row, 8(e)) -- one process, ncclCommInitAll, one host
thread per GPU.  A one-device group runs everywhere; the two-device cases need a box with two B200s (gpurun --gpus 2)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _ndev():
    from sparsesolvers._lib import lib
    return lib().ssb200_device_count()


@pytest.fixture(scope="module")
def case():
    import sparsesolvers as ss
    from sparsesolvers import synth
    A = synth.dictionary(128, 512, dtype=np.float32, seed=41)
    Y, _ = synth.sparse_signals(A, 301, 9, seed=42)             # odd count: shards of unequal length
    X, it, err = ss.Homotopy(A).solve_batch(Y, 1e-3, 40)
    return ss, A, Y, X, it, err


@pytest.mark.parametrize("ndev", [1, 2])
def test_group_host_entry_equals_single_handle(case, ndev):
    ss, A, Y, X, it, err = case
    if _ndev() < ndev:
        pytest.skip(f"needs {ndev} GPUs")
    g = ss.HomotopyGroup(A, devices=list(range(ndev)))
    assert len(g) == ndev
    Xg, itg, eg = g.solve_batch(Y, 1e-3, 40)
    assert (itg == it).all()
    np.testing.assert_array_equal(Xg, X)          # a signal's arithmetic does not depend on the shard it travels in
    np.testing.assert_array_equal(eg, err)


@pytest.mark.parametrize("ndev", [1, 2])
def test_group_device_entry_scatter_and_gather(case, ndev):
    import torch
    ss, A, Y, X, it, err = case
    if _ndev() < ndev:
        pytest.skip(f"needs {ndev} GPUs")
    g = ss.HomotopyGroup(np.asfortranarray(A), devices=list(range(ndev)))     # a column-major view of the same dictionary
    dX, dit, derr = g.solve_batch(torch.from_numpy(Y).cuda(0), 1e-3, 40)
    assert dX.device.index == 0
    assert (dit.cpu().numpy().astype(np.uint32) == it).all()
    np.testing.assert_array_equal(dX.cpu().numpy(), X)
    np.testing.assert_array_equal(derr.cpu().numpy(), err)


def test_group_argument_checks(case):
    ss, A, Y, X, it, err = case
    with pytest.raises(ss.SolverError):
        ss.HomotopyGroup(A, devices=[0, 0])                   # a device may appear once
    with pytest.raises(ss.SolverError, match="Compute Mode Unavailable"):
        ss.HomotopyGroup(A, devices=[99])
    g = ss.HomotopyGroup(A, devices=[0])
    with pytest.raises(TypeError):
        g.solve_batch(Y.astype(np.float64))
    Xe, ite, ee = g.solve_batch(np.zeros((0, 128), dtype=np.float32), 1e-3, 10)
    assert Xe.shape == (0, 512)
