"""The residual-correlation contraction Q = A^T P alone: both kernels against float64 numpy."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ss():
    import sparsesolvers
    return sparsesolvers


def _rel_err(Q, A, P):
    ref = P.astype(np.float64) @ A.astype(np.float64)
    scale = np.abs(P.astype(np.float64)) @ np.abs(A.astype(np.float64))   # sum |a||p|: the rounding scale
    return (np.abs(Q - ref) / scale).max()


@pytest.mark.parametrize("m,n,rows", [(512, 2048, 1), (512, 2048, 3), (512, 2048, 300), (504, 1216, 77), (96, 160, 40)])
def test_simt_f32(ss, m, n, rows):
    rng = np.random.default_rng(m + rows)
    A = (rng.standard_normal((m, n)) / np.sqrt(m)).astype(np.float32)
    P = rng.standard_normal((rows, m)).astype(np.float32)
    Q, _ = ss.Homotopy(A).debug_contract(P, "simt")
    assert _rel_err(Q, A, P) < 5e-7


@pytest.mark.parametrize("path", ["tcgen05", "tcgen05-tf32"])
@pytest.mark.parametrize("m,n,rows", [(512, 2048, 256), (512, 2048, 1000), (504, 1216, 77), (96, 160, 40), (4096, 1024, 513),
                                      (512, 2048, 20)])
def test_tcgen05_3split(ss, m, n, rows, path):
    """The three-product split must stay at float32 level in both containers (fp16 with power-of-two row scales, tf32):
    the error is measured against sum |a||p| and held to a few fp32 ulps.  Row scales span 12 decades and the atoms
    5, so the fp16 range only works if every row really is rescaled."""
    rng = np.random.default_rng(m + rows)
    A = (rng.standard_normal((m, n)) / np.sqrt(m) * np.exp(rng.uniform(-6, 6, size=(1, n)))).astype(np.float32)
    P = (rng.standard_normal((rows, m)) * np.exp(rng.uniform(-14, 14, size=(rows, 1)))).astype(np.float32)
    s = ss.Homotopy(A)
    Q, _ = s.debug_contract(P, path)
    Qs, _ = s.debug_contract(P, "simt")
    assert np.isfinite(Q).all()
    e_tc, e_simt = _rel_err(Q, A, P), _rel_err(Qs, A, P)
    print(f"rel err {path} {e_tc:.3g} simt {e_simt:.3g}")
    assert e_tc < 5e-6     # tensor-core accumulation rounds toward zero: grows with M


def test_f16_containers_match_tf32(ss):
    """Same significand width, same products: the two containers agree to accumulation-order noise."""
    rng = np.random.default_rng(3)
    A = (rng.standard_normal((512, 2048)) / np.sqrt(512)).astype(np.float32)
    P = rng.standard_normal((300, 512)).astype(np.float32)
    s = ss.Homotopy(A)
    Qh, _ = s.debug_contract(P, "tcgen05")
    Qt, _ = s.debug_contract(P, "tcgen05-tf32")
    scale = np.abs(P.astype(np.float64)) @ np.abs(A.astype(np.float64))
    assert (np.abs(Qh - Qt) / scale).max() < 1e-6


@pytest.mark.parametrize("m,n,rows", [(200, 300, 50), (512, 2048, 4096), (504, 1216, 77), (96, 160, 5), (4096, 1024, 130)])
def test_f64(ss, m, n, rows):
    """float64 batches: the DMMA kernel (mma.sync.m8n8k4.f64, the default for more than 4 signals) and the plain DFMA tiles it
    replaced, both against numpy; error measured against sum |a||p| as above."""
    rng = np.random.default_rng(m + rows)
    A = rng.standard_normal((m, n)) / np.sqrt(m)
    P = rng.standard_normal((rows, m)) * np.exp(rng.uniform(-20, 20, size=(rows, 1)))
    s = ss.Homotopy(A)
    Q, ms_dmma = s.debug_contract(P, "simt", reps=3)
    Qf, ms_dfma = s.debug_contract(P, "fma", reps=3)
    e, ef = _rel_err(Q, A, P), _rel_err(Qf, A, P)
    print(f"f64 {m}x{n}x{rows}: DMMA rel err {e:.2e} ({ms_dmma:.3f} ms, {2e-9 * m * n * rows / ms_dmma:.1f} TFLOP/s)  "
          f"DFMA {ef:.2e} ({ms_dfma:.3f} ms, {2e-9 * m * n * rows / ms_dfma:.1f} TFLOP/s)")
    assert e < 1e-15 and ef < 1e-15
    assert np.abs(Q - Qf).max() <= 4e-16 * (np.abs(P) @ np.abs(A)).max()


def test_fp64_peak_probe(ss):
    dfma, dmma = ss.fp64_peak("dfma"), ss.fp64_peak("dmma")
    print(f"FP64 peak probes: DFMA {dfma:.1f} TFLOP/s, DMMA {dmma:.1f} TFLOP/s")
    assert 5 < dfma < 200 and 5 < dmma < 400
