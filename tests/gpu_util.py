"""Helpers shared by the -m gpu parity tests."""
import os

import numpy as np

import parity


def c0_inf(A, Y):
    return np.abs(np.asarray(Y, dtype=np.float64) @ np.asarray(A, dtype=np.float64)).max(axis=1)


def have_ref():
    from oracle import ref
    return ref.available() and os.path.exists("/root/repo/oracle/_ref/libcblas_fwd.so")


_MEMO = {}


def _key(who, A, Y, tol, max_iter):
    import hashlib
    return (who, A.shape, str(A.dtype), hashlib.md5(np.ascontiguousarray(A)).hexdigest(), hashlib.md5(np.ascontiguousarray(Y)).hexdigest(),
            float(tol), int(max_iter))


def _cpu(who, A, Y, tol, max_iter):
    """memoised across test parametrisations (the CPU side is the slow one)"""
    from oracle import port, ref
    k = _key(who, A, Y, tol, max_iter)
    if k not in _MEMO:
        if who == "ref":
            ref.set_blas_threads(1)
            _MEMO[k] = ref.homotopy(A, Y, tol, max_iter, threads=len(os.sched_getaffinity(0)))
        else:
            _MEMO[k] = port.homotopy(A, Y, tol, max_iter)
    return _MEMO[k]


def cpu_reference(A, Y, tol, max_iter):
    """(X, iters, errs) of the reference's CPU path on these inputs: the UNMODIFIED reference build (oracle/_ref, present
    on the GPU box) when it is there, else the plain-C restatement.  Second value: which one."""
    who = "ref" if have_ref() else "port"
    return _cpu(who, A, Y, tol, max_iter), who


def cpu_floor(A, Y, tol, max_iter, primary):
    """Mismatch counts of the two CPU builds against each other on these inputs (float32 noise floor of the gate);
    None when only one of them is at hand."""
    if not have_ref() or np.dtype(A.dtype) != np.dtype(np.float32):
        return None
    Xp, itp, ep = _cpu("port", A, Y, tol, max_iter)
    return parity.counts(parity.compare(Xp, itp, ep, primary[0], primary[1], primary[2], c0_inf(A, Y)))


def check_against_cpu(solver, A, Y, tol, max_iter, what, noise_rate=0.02):
    """The parity gate against the reference build (the restatement where that is absent), held to the floor the two CPU
    builds show against each other on the same inputs.  Returns (counts, floor, (X, iters, errs), (Xr, itr, er))."""
    primary, who = cpu_reference(A, Y, tol, max_iter)
    floor = cpu_floor(A, Y, tol, max_iter, primary)
    X, it, err = solver.solve_batch(Y, tol, max_iter)
    r = parity.compare(X, it, err, primary[0], primary[1], primary[2], c0_inf(A, Y))
    c = parity.assert_parity(r, A.dtype, f"{what} [vs {who}]", noise_rate=noise_rate, floor=floor)
    print(f"PARITY {what} vs {who}: {c} floor(port vs ref): {floor}")
    return c, floor, (X, it, err), primary


def check_homotopy(solver, A, Y, tol, max_iter, Xr, itr, er, what, noise_rate=0.02):
    X, it, err = solver.solve_batch(Y, tol, max_iter)
    r = parity.compare(X, it, err, Xr, itr, er, c0_inf(A, Y))
    return parity.assert_parity(r, A.dtype, what, noise_rate=noise_rate), (X, it, err)


def _next_perm(v):
    v = list(v)
    i = len(v) - 2
    while i >= 0 and v[i] >= v[i + 1]:
        i -= 1
    if i < 0:
        return v[::-1]
    j = len(v) - 1
    while v[j] <= v[i]:
        j -= 1
    v[i], v[j] = v[j], v[i]
    v[i + 1:] = reversed(v[i + 1:])
    return v


def _permute(v, n):
    for _ in range(n):
        v = _next_perm(v)
    return v


def permutation_case(M, N, dtype, noise, skip, seed=0):
    """The reference's permutations_test inputs (src/solvers/test_util.h:204-257): column n of A is the
    (n*skip)-th next_permutation of 1..M plus N(0, noise) sensing noise; signal n is the same permutation
    plus N(0, noise) signal noise.  The reference draws from xtensor's RNG; numpy's is used here, so the
    expected answers are taken from the CPU oracle on the same draws, not assumed."""
    rng = np.random.default_rng(seed)
    col = _permute(list(range(1, M + 1)), skip)
    A = rng.normal(0, noise, size=(M, N))
    c = list(col)
    for n in range(N):
        A[:, n] += c
        c = _permute(c, skip)
    sigs, c = [], list(col)
    for n in range(N):
        sigs.append(np.array(c, dtype=np.float64) + rng.normal(0, noise, size=M))
        c = _permute(c, skip)
    return A.astype(dtype), np.array(sigs).astype(dtype)
