"""GPU parity tests of the homotopy path, through the C ABI (sparsesolvers -> libssb200.so):
  - the committed golden vectors (outputs of the unmodified reference, tests/golden/make_golden.py)
  - the CPU restatement on seeded inputs, float32 and float64, both contraction kernels
  - the reference's behavioural tests (src/solvers/test_util.h, bindings/python/tests/test_binding.py)
  - edge cases and size-independent properties at larger sizes
"""
import glob
import itertools
import os

import numpy as np
import pytest

from conftest import GOLDEN
import parity
from gpu_util import c0_inf, check_homotopy, check_against_cpu, cpu_reference, cpu_floor, have_ref

pytestmark = pytest.mark.gpu

HOMOTOPY_GOLDEN = sorted(os.path.basename(p) for p in glob.glob(os.path.join(GOLDEN, "homotopy_*.npz")))


@pytest.fixture(scope="module")
def ss():
    import sparsesolvers
    return sparsesolvers


@pytest.mark.parametrize("path", ["simt", "auto", "tcgen05-tf32"])
@pytest.mark.parametrize("name", HOMOTOPY_GOLDEN)
def test_golden(ss, name, path):
    g = np.load(os.path.join(GOLDEN, name))
    A = g["A"]
    if "colmajor" in name:
        A = np.asfortranarray(A)
    s = ss.Homotopy(A)
    s.set_contraction_path(path)
    check_homotopy(s, g["A"], g["Y"], float(g["tol"]), int(g["max_iter"]), g["X"], g["iters"], g["errs"], name)


_CPU_CACHE = {}


def _cpu_case(dtype, m, n, k, batch):
    """inputs + the CPU results (reference build where present, else the restatement) + the CPU-against-CPU floor, once per shape"""
    key = (np.dtype(dtype).name, m, n, k, batch)
    if key not in _CPU_CACHE:
        from sparsesolvers import synth
        A = synth.dictionary(m, n, dtype=dtype, seed=5 + m)
        Y, _ = synth.sparse_signals(A, batch, k, seed=6 + n)
        mi = 2 * k + 16
        primary, who = cpu_reference(A, Y, 1e-3, mi)
        _CPU_CACHE[key] = (A, Y, mi, primary, who, cpu_floor(A, Y, 1e-3, mi, primary))
    return _CPU_CACHE[key]


@pytest.mark.parametrize("mode", ["auto", "exact"])
@pytest.mark.parametrize("path", ["simt", "auto", "tcgen05-tf32"])
@pytest.mark.parametrize("dtype,m,n,k,batch", [(np.float32, 128, 512, 10, 300), (np.float64, 128, 512, 10, 200),
                                               (np.float32, 256, 1024, 16, 256), (np.float32, 504, 1216, 20, 130),
                                               (np.float64, 100, 230, 7, 70), (np.float32, 512, 2048, 32, 192),
                                               (np.float64, 512, 2048, 32, 64)])
def test_vs_cpu(ss, dtype, m, n, k, batch, path, mode):
    """Seeded inputs against the reference's CPU path: the unmodified reference build (oracle/_ref) where it is present --
    it is on the GPU box -- else the plain-C restatement; every contraction kernel, both residual modes.  float64: exact
    iteration counts and supports, 1e-10.  float32: the same gate held to the floor the two CPU builds show against each
    other on these very inputs (parity.assert_parity), nothing above 1e-3."""
    A, Y, mi, primary, who, floor = _cpu_case(dtype, m, n, k, batch)
    s = ss.Homotopy(A)
    s.set_contraction_path(path)
    s.set_residual_mode(mode)
    X, it, err = s.solve_batch(Y, 1e-3, mi)
    r = parity.compare(X, it, err, primary[0], primary[1], primary[2], c0_inf(A, Y))
    what = f"{np.dtype(dtype).name} {m}x{n} k={k} [{path}, {mode} -> {s.residual_info()[0]}] vs {who}"
    c = parity.assert_parity(r, A.dtype, what, floor=floor)
    print("PARITY", what, c, "floor(port vs ref):", floor, s.counters())


@pytest.mark.parametrize("name", HOMOTOPY_GOLDEN)
def test_golden_one_call_per_signal(ss, name):
    """The single-signal entry (one cooperative kernel per solve, homotopy_single.cuh) against the reference's outputs,
    quirk set included; 48 signals per fixture."""
    g = np.load(os.path.join(GOLDEN, name))
    A = np.asfortranarray(g["A"]) if "colmajor" in name else g["A"]
    s = ss.Homotopy(A)
    n = min(48, len(g["Y"]))
    X = np.empty((n, A.shape[1]), dtype=A.dtype); it = np.empty(n, dtype=np.uint32); err = np.empty(n)
    for b in range(n):
        X[b], rep = s.solve(g["Y"][b], float(g["tol"]), int(g["max_iter"]))
        it[b], err[b] = rep.iter, rep.solution_error
    assert s.counters()["launches"] <= 4 + n, "the single-signal kernel did not run"   # dictionary set-up (pack / Gram / split / coherence) + one kernel per solve
    r = parity.compare(X, it, err, g["X"][:n], g["iters"][:n], g["errs"][:n], c0_inf(g["A"], g["Y"][:n]))
    print(parity.assert_parity(r, A.dtype, name + " [single]"))


def test_single_solve_equals_batch_row(ss):
    from sparsesolvers import synth
    A = synth.dictionary(64, 256, dtype=np.float32, seed=1)
    Y, _ = synth.sparse_signals(A, 5, 4, seed=2)
    s = ss.Homotopy(A)
    Xb, itb, eb = s.solve_batch(Y, 1e-3, 30)
    for b in range(5):
        x, rep = s.solve(Y[b], 1e-3, 30)
        assert rep.iter == itb[b]
        np.testing.assert_allclose(x, Xb[b], rtol=2e-4, atol=1e-6)


@pytest.mark.parametrize("dtype,shape,k,max_iter", [(np.float32, (256, 1024), 16, 48), (np.float64, (128, 512), 8, 30), (np.float32, (512, 2048), 32, 80)])
@pytest.mark.parametrize("mode", ["incremental", "exact"])
def test_single_solve_cluster_form_equals_the_grid_form(ss, dtype, shape, k, max_iter, mode, monkeypatch):
    """homotopy_single_kernel as one thread-block cluster per signal (barrier.cluster, exchange through distributed shared memory)
    against the cooperative-grid form (grid.sync, exchange through global scratch): same arithmetic in the same order, so the
    same bits -- coefficients, iteration counts and the reported error; also at cluster size 8 (the portable size)."""
    from sparsesolvers import synth
    A = synth.dictionary(*shape, dtype=dtype, seed=31)
    Y, _ = synth.sparse_signals(A, 6, k, seed=32)
    Y[5] = np.random.default_rng(33).standard_normal(shape[0]).astype(dtype)      # a signal that runs to max_iterations
    s = ss.Homotopy(A)
    s.set_residual_mode(mode)
    got = {}
    for form, env in (("grid", {"SSB200_SINGLE_CLUSTER": "0"}), ("cluster16", {"SSB200_SINGLE_CLUSTER": "1"}), ("cluster8", {"SSB200_SINGLE_CLUSTER": "1", "SSB200_CLUSTER_SIZE": "8"})):
        monkeypatch.delenv("SSB200_CLUSTER_SIZE", raising=False)
        for key, v in env.items():
            monkeypatch.setenv(key, v)
        got[form] = [s.solve(Y[b], 1e-3, max_iter) for b in range(len(Y))]
    for form in ("cluster16", "cluster8"):
        for (xg, rg), (xc, rc) in zip(got["grid"], got[form]):
            assert rc.iter == rg.iter and rc.solution_error == rg.solution_error, (form, rc.iter, rg.iter)
            np.testing.assert_array_equal(xc, xg)
    assert max(r.iter for _, r in got["grid"]) == max_iter


def test_single_solve_out_of_slots_falls_back_to_the_lockstep_path(ss):
    """The single-signal kernel is launched with min(max_iterations + 1, N, M + 1) active-set slots (a full-rank active set has at
    most M atoms; the reference's bench passes max_iterations = N).  A solve that asks for more -- here: tolerance ~0, so it keeps
    adding atoms past K = M on rounding residue -- reports so, and the call is answered by the lockstep path instead: the same
    result as that path gives for the signal inside a batch."""
    rng = np.random.default_rng(5)
    A = rng.standard_normal((8, 64))
    A /= np.linalg.norm(A, axis=0)
    y = rng.standard_normal(8)
    s = ss.Homotopy(A)
    s.reset_counters()
    x, rep = s.solve(y, 1e-300, 40)
    assert s.counters()["launches"] > 3, "the fall-back did not run"
    Xb, itb, eb = s.solve_batch(np.stack([y, y]), 1e-300, 40)
    assert rep.iter == itb[0] and rep.iter > 9
    np.testing.assert_array_equal(x, Xb[0])
    # and a solve that stays within M + 1 slots is served by the single kernel: one launch
    s.reset_counters()
    x2, rep2 = s.solve(y, 1e-3, 40)
    assert s.counters()["launches"] == 1 and rep2.iter <= 12


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_smoke_identity(ss, dtype):
    """test_util.h:27-55, test_binding.py:9-20: x == signal exactly, iter == 1, solution_error == 0."""
    N = 5
    s = ss.Homotopy(np.eye(N, dtype=dtype))
    for n in range(N):
        y = np.zeros(N, dtype=dtype)
        y[n] = 1
        x, rep = s.solve(y, 1e-3, N)
        assert (x == y).all() and rep.iter == 1 and rep.solution_error == 0.0


def test_binding_defaults(ss):
    """test_binding.py:9-20 calls solve(signal) with the default tolerance 10*eps and 100 iterations."""
    for dtype in (np.float32, np.float64):
        s = ss.Homotopy(np.identity(10, dtype=dtype))
        y = np.zeros(10, dtype=dtype)
        y[3] = 1
        x, rep = s.solve(y)
        assert (x == y).all() and rep.iter == 1 and rep.solution_error == 0


def test_strided_views(ss):
    """test_util.h:57-92 and test_binding.py:31-68: column-subset, row-subset and transposed views."""
    buf = np.zeros((5, 10), dtype=np.float32)
    buf[:, 5:] = np.eye(5, dtype=np.float32)
    s = ss.Homotopy(buf[:, 5:])
    for n in range(5):
        y = np.zeros(5, dtype=np.float32)
        y[n] = 1
        x, _ = s.solve(y, 1e-3, 5)
        assert (x == y).all()
    big = np.zeros((20, 10), dtype=np.float64)
    big[::2, :] = np.eye(10)
    s = ss.Homotopy(big[::2, :])
    y = np.zeros(10)
    y[7] = 1
    x, rep = s.solve(y)
    assert (x == y).all() and rep.iter == 1
    T = np.triu(np.ones((6, 6), dtype=np.float32)) / 6
    st, sc = ss.Homotopy(T.T), ss.Homotopy(np.ascontiguousarray(T.T))
    y = T.T[:, 2].copy()
    xa, ra = st.solve(y, 1e-3, 12)
    xb, rb = sc.solve(y, 1e-3, 12)
    assert ra.iter == rb.iter and np.array_equal(xa, xb)
    # a strided signal view
    y2 = np.zeros(20)
    y2[::2][7] = 1
    x, rep = s.solve(y2[::2])
    assert x[7] == 1 and rep.iter == 1


def test_noisy_signal(ss):
    """test_util.h:94-126"""
    rng = np.random.default_rng(0)
    N = 50
    s = ss.Homotopy(np.eye(N, dtype=np.float32))
    for n in range(0, N, 7):
        y = rng.uniform(0, 0.01, size=N).astype(np.float32)
        y[n] += 1
        x, rep = s.solve(y, 0.01, N)
        assert (x > 0.01).sum() == 1 and np.argmax(x) == n
        assert 1 <= rep.iter <= N and (rep.iter == N or rep.solution_error <= 0.01)


@pytest.mark.parametrize("shape,dtype", [((10, 10), np.float32), ((25, 10), np.float32), ((10, 25), np.float32),
                                         ((10, 25), np.float64)])
def test_permutations(ss, shape, dtype):
    """test_util.h:204-257 (homotopy_test.cpp:48-61)"""
    M, N = shape
    rng = np.random.default_rng(0)
    perms = itertools.permutations(range(1, M + 1))
    A = np.stack([np.array(next(perms), dtype=np.float64) for _ in range(N)], axis=1)
    A /= np.abs(A).sum(axis=0, keepdims=True)
    A = A.astype(dtype)
    s = ss.Homotopy(A)
    Y = (A.T.astype(np.float64) + rng.standard_normal((N, M)) * 1e-4).astype(dtype)
    X, it, err = s.solve_batch(Y, 1e-3, N)
    assert (np.argmax(X, axis=1) == np.arange(N)).all()


@pytest.mark.parametrize("shape", [(100, 25), (25, 100)])
def test_noisy_patterns(ss, shape):
    """test_util.h:136-197: a patterned needle in an l1-normalised Gaussian haystack; reconstruction within 5*ERROR."""
    M, N = shape
    ERROR = 0.05
    rng = np.random.default_rng(0)
    for n in range(0, N, max(1, N // 5)):
        A = rng.normal(0.5, 0.1, size=(M, N)).astype(np.float32)
        A[::2, n] += 0.5
        ss.norm_l1(A)
        np.testing.assert_allclose(np.abs(A).sum(axis=0), 1, rtol=1e-5)
        y = A[:, n].copy()
        s = ss.Homotopy(A)
        x, rep = s.solve(y, ERROR / 10, N)
        assert np.argmax(x) == n
        yr = ss.reconstruct_signal(A, x)
        assert np.abs(yr - y).max() < 5 * ERROR


def test_quirk_sign_init(ss):
    """homotopy-cpu.cpp:223-227: d0 = sign(|c|) -- un-normalised signals follow the reference's own path."""
    g = np.load(os.path.join(GOLDEN, "homotopy_f64_quirk_64x256_k8.npz"))
    s = ss.Homotopy(g["A"])
    X, it, err = s.solve_batch(g["Y"], float(g["tol"]), int(g["max_iter"]))
    assert (it == g["iters"]).all()
    np.testing.assert_allclose(X, g["X"], rtol=1e-8, atol=1e-10 * np.abs(g["X"]).max())


def test_edge_cases(ss):
    from oracle import port
    from sparsesolvers import synth
    A = synth.dictionary(48, 100, dtype=np.float32, seed=8)     # N, M not multiples of the tile sizes
    s = ss.Homotopy(A)
    with pytest.raises(ValueError):
        s.solve_batch(np.zeros((3, 48), dtype=np.float32), 1e-3, 10, out=np.zeros((3, 100), dtype=np.float64))   # wrong dtype
    with pytest.raises(ValueError):
        s.solve_batch(np.zeros((3, 48), dtype=np.float32), 1e-3, 10, out=np.zeros((2, 100), dtype=np.float32))   # too short
    with pytest.raises(ValueError):
        s.solve_batch(np.zeros((3, 48), dtype=np.float32), 1e-3, 10, out=np.zeros((3, 200), dtype=np.float32)[:, ::2])
    X, it, err = s.solve_batch(np.zeros((0, 48), dtype=np.float32), 1e-3, 10)      # empty batch
    assert X.shape == (0, 100) and it.shape == (0,)
    Y, _ = synth.sparse_signals(A, 7, 3, seed=9)
    Y[3] = 0                                                     # an all-zero signal
    (Xr, itr, er), _ = cpu_reference(A, Y, 1e-3, 20)
    Xg, itg, eg = s.solve_batch(Y, 1e-3, 20)
    assert (itg == itr).all()
    assert itg[3] == 1 and not Xg[3].any()
    # max_iter = 1 and max_iter > N
    for mi in (1, 150):
        Xr, itr, er = port.homotopy(A, Y[:3], 1e-3, mi)
        Xg, itg, eg = s.solve_batch(Y[:3], 1e-3, mi)
        assert (itg == itr).all()
        np.testing.assert_allclose(Xg, Xr, rtol=1e-3, atol=1e-5)
    with pytest.raises(TypeError):
        s.solve(Y[0].astype(np.float64))                         # noconvert, binding.cpp:108
    with pytest.raises(ValueError):
        s.solve(np.zeros(5, dtype=np.float32))


def test_refresh(ss):
    """the reference's non-owning view lets callers edit A between solves (homotopy_bench.cpp:38-48)."""
    A = np.eye(6, dtype=np.float32)
    s = ss.Homotopy(A)
    y = np.zeros(6, dtype=np.float32); y[2] = 1
    A2 = 2 * A
    s.refresh(A2)
    x, rep = s.solve(y, 1e-3, 6)
    assert x[2] == 0.5 and rep.iter == 1


def test_full_size_properties(ss):
    """C2 shape (512x2048, k=32) on a batch too large for the CPU oracle in test time: size-independent properties.
    Converged signals reconstruct y, recover the planted support, and take k..k+4 iterations."""
    from sparsesolvers import synth
    A = synth.dictionary(512, 2048, dtype=np.float32, seed=1234 + 2)
    Y, X0 = synth.sparse_signals(A, 4096, 32, seed=99)
    s = ss.Homotopy(A)
    X, it, err = s.solve_batch(Y, 1e-3, 80)
    conv = it < 80
    assert conv.mean() > 0.995
    assert (err[conv] <= 1e-3).all()
    assert it[conv].min() >= 32 and it[conv].max() <= 40
    sup = parity.support(X, 1e-5)
    assert ((sup == (X0 != 0)).all(axis=1))[conv].mean() > 0.995
    R = Y - s.reconstruct(X)
    assert np.abs(R[conv]).max() < 5e-3
    # idempotence of the batch split: the same signals in a different batch composition give the same answers
    # (stragglers -- signals still running six iterations after the chunk has thinned out -- are finished by the cooperative
    # kernel from wherever their chunk happened to be: they agree to rounding; everything else bit for bit)
    X2, it2, _ = s.solve_batch(Y[1000:1100], 1e-3, 80)
    usual = it[1000:1100] <= 42
    assert usual.mean() > 0.95
    assert (it2[usual] == it[1000:1100][usual]).all()
    np.testing.assert_array_equal(X2[usual], X[1000:1100][usual])
    np.testing.assert_allclose(X2[~usual], X[1000:1100][~usual], rtol=0, atol=2e-2 * np.abs(X).max())


@pytest.mark.parametrize("mode", ["auto", "exact"])
def test_c3_face_recognition_shape(ss, mode):
    """BASELINE config 3: 504x1216 unit-l2 dictionary, 10 % gross corruption, max_iter 100 (binding default).
    Nothing converges: every signal runs all 100 iterations and the active set grows to ~100.
    The late path is rounding-chaotic (the two CPU builds themselves drift apart between 60 and 100 iterations), so both
    lengths are held to the floor those two builds show on these very inputs."""
    from sparsesolvers import synth
    A, Y, labels = synth.face_recognition(batch=96, seed=77)
    s = ss.Homotopy(A)
    s.set_residual_mode(mode)
    check_against_cpu(s, A, Y, 1e-3, 60, f"C3 @60 [{mode}]")
    primary, who = cpu_reference(A, Y, 1e-3, 100)
    floor = cpu_floor(A, Y, 1e-3, 100, primary)
    X, it, err = s.solve_batch(Y, 1e-3, 100)
    Xr, itr, er = primary
    assert (it == itr).all() and (it == 100).all()
    c = parity.counts(parity.compare(X, it, err, Xr, itr, er, c0_inf(A, Y)))
    print(f"PARITY C3 @100 [{mode}] vs {who}:", c, "floor(port vs ref):", floor)
    if floor is not None:
        assert c["support"] <= parity.allowance(floor["support"]) and c["coef"] <= parity.allowance(floor["coef"])
    else:
        assert c["support"] <= 0.06 * len(it) and c["coef"] <= 0.10 * len(it)
    per = 1216 // 38
    cls = np.abs(X).reshape(len(X), 38, per).sum(axis=2).argmax(axis=1)
    cls_ref = np.abs(Xr).reshape(len(X), 38, per).sum(axis=2).argmax(axis=1)
    assert (cls == cls_ref).all() and (cls == labels).mean() > 0.95


@pytest.mark.parametrize("mode", ["auto", "exact"])
def test_c5_large_dictionary_shape(ss, mode):
    """BASELINE config 5 shape: 4096x16384, k=128, max_iter 272 -- wide-dictionary step-kernel variant (c, q re-read),
    inverse too large for shared memory at the end of the path; a few signals so the CPU side finishes."""
    from sparsesolvers import synth
    A = synth.dictionary(4096, 16384, dtype=np.float32, seed=1234 + 5)
    Y, X0 = synth.sparse_signals(A, 24, 128, seed=55)
    s = ss.Homotopy(A)
    s.set_residual_mode(mode)
    c, floor, (X, it, err), _ = check_against_cpu(s, A, Y, 1e-3, 272, f"C5 [{mode}]", noise_rate=0.1)
    assert ((parity.support(X, 1e-5) == (X0 != 0)).all(axis=1)).mean() > 0.9


@pytest.mark.parametrize("dtype,max_iter", [(np.float32, 500), (np.float32, 1023), (np.float64, 1023)])
def test_large_max_iterations(ss, dtype, max_iter):
    """The reference takes any max_iterations (its bench passes N, homotopy_bench.cpp:41).  The active-set capacity is
    min(max_iterations + 1, N): above 384 slots the step kernel runs 256 threads, above 768 the 16-slots-per-thread
    variant with the inverse worked on in HBM.  Converging signals must come out as the oracle has them."""
    from oracle import port
    from sparsesolvers import synth
    A = synth.dictionary(128, 1024, dtype=dtype, seed=77)
    Y, _ = synth.sparse_signals(A, 48, 8, seed=78)
    (Xr, itr, er), _ = cpu_reference(A, Y, 1e-3, max_iter)
    assert itr.max() < 40
    s = ss.Homotopy(A)
    stats, _ = check_homotopy(s, A, Y, 1e-3, max_iter, Xr, itr, er, f"{np.dtype(dtype).name} max_iter={max_iter}")
    x, rep = s.solve(Y[0], 1e-3, max_iter)              # single-signal entry with the same capacity
    assert rep.iter == itr[0]
    np.testing.assert_allclose(x, Xr[0], rtol=0, atol=1e-4 * np.abs(Xr[0]).max())


def test_long_runs_large_active_set(ss):
    """Signals the reference walks to max_iterations (sign quirk, homotopy-cpu.cpp:223-224) with a large budget: 900
    iterations on the 16-slot variant.  The trajectory is chaotic at that length, so the check is the report the
    reference gives (iter == max_iterations for the same signals, finite coefficients), not the digits."""
    from oracle import port
    from sparsesolvers import synth
    A = synth.dictionary(96, 1024, dtype=np.float32, seed=81)
    Y, _ = synth.sparse_signals(A, 24, 6, seed=82, normalise_sign=False)
    (Xr, itr, er), _ = cpu_reference(A, Y, 1e-3, 900)
    X, it, err = ss.Homotopy(A).solve_batch(Y, 1e-3, 900)
    assert np.isfinite(X).all()
    long_ref, short_ref = itr == 900, itr <= 10
    assert long_ref.any() and short_ref.any()
    # converging signals agree exactly; of those the reference walks to the budget, most must do so here as well
    assert (it[short_ref] == itr[short_ref]).all()
    assert (it[long_ref] == 900).sum() * 2 >= long_ref.sum(), (it[long_ref], itr[long_ref])
    assert (it[~short_ref] > 10).all()


@pytest.mark.parametrize("M,N,solves", [(16, 16, 8), (16, 64, 8), (64, 64, 8), (64, 256, 16), (256, 256, 16), (256, 1024, 16), (512, 1024, 12)])
def test_reference_bench_grid_cells(ss, M, N, solves):
    """The reference's own benchmark workload (src/solvers/homotopy_bench.cpp:14-48; bench.py --config GRID): a +1 pattern
    planted in column n of a N(0.5, 0.1) haystack, edited IN PLACE between solves (the handle follows through refresh),
    tolerance 0.1, max_iterations = N.  Every atom is nearly parallel to every other (mutual coherence ~0.97), so the handle
    takes the exact-residual mode by itself.  The needle must come out on top; iteration counts are held to what the two CPU
    builds of the reference show against each other on the same solves: at least as many equal counts as the restatement has
    (less the allowance), and as many solves running to max_iterations as the reference build, +-2 beyond the restatement's own gap."""
    import bench
    from oracle import port
    hay, sig = bench.grid_workload(M, N)
    s = ss.Homotopy(hay)
    it_g, it_r, it_p = (np.zeros(solves, dtype=np.int64) for _ in range(3))
    for i in range(solves):
        n = i % N
        hay[::2, n] += 1.0
        s.refresh(hay)
        assert s.residual_info()[0] == "exact" and s.residual_info()[1] > 0.9
        x, rep = s.solve(sig, 0.1, N)
        (xr, itr, er), who = cpu_reference(hay, sig[None, :], 0.1, N)
        xp, itp, ep = port.homotopy(hay, sig[None, :], 0.1, N)
        hay[::2, n] -= 1.0
        # (on the larger cells the reference itself breaks down on some solves -- its inverse turns singular and x overflows,
        # e.g. 512x1024 solve 0: 888 iterations, solution_error NaN, max|x| 2.5e38 -- so the needle is asked for where it finds it)
        if np.isfinite(xr[0]).all() and int(np.argmax(np.abs(xr[0]))) == n and er[0] <= 0.1:
            assert int(np.argmax(np.abs(x))) == n
        assert rep.solution_error <= 0.1 or rep.iter == N or not np.isfinite(rep.solution_error)
        it_g[i], it_r[i], it_p[i] = rep.iter, itr[0], itp[0]
        if itr[0] <= 8 and itp[0] == itr[0]:
            assert rep.iter == itr[0]
            np.testing.assert_allclose(x, xr[0], rtol=0, atol=1e-3 * np.abs(xr[0]).max())
    eq_g, eq_p = int((it_g == it_r).sum()), int((it_p == it_r).sum())
    hm_g, hm_r, hm_p = int((it_g == N).sum()), int((it_r == N).sum()), int((it_p == N).sum())
    print(f"GRID {M}x{N}: iters equal to {who}: gpu {eq_g}/{solves}, restatement {eq_p}/{solves}; max_iter hits gpu {hm_g} ref {hm_r} restatement {hm_p}; "
          f"mean iters gpu {it_g.mean():.1f} ref {it_r.mean():.1f} restatement {it_p.mean():.1f}")
    assert eq_g >= eq_p - parity.allowance(solves - eq_p) // 2, f"{eq_g}/{solves} iteration counts equal to the reference's (restatement: {eq_p})"
    assert abs(hm_g - hm_r) <= abs(hm_p - hm_r) + 2, f"max_iter hits: gpu {hm_g}, reference {hm_r}, restatement {hm_p}"


def test_residual_mode_selection(ss):
    """auto = exact on coherent dictionaries (the reference bench's haystack), incremental on incoherent ones; both can be forced."""
    import bench
    from sparsesolvers import synth
    hay, _ = bench.grid_workload(64, 256)
    s = ss.Homotopy(hay)
    mode, mu = s.residual_info()
    assert mode == "exact" and mu > 0.9
    s.set_residual_mode("incremental")
    assert s.residual_info()[0] == "incremental"
    g = ss.Homotopy(synth.dictionary(256, 1024, dtype=np.float32, seed=3))
    mode, mu = g.residual_info()
    assert mode == "incremental" and 0.05 < mu < 0.6
    g.set_residual_mode("exact")
    assert g.residual_info()[0] == "exact"
    with pytest.raises(KeyError):
        g.set_residual_mode("sometimes")


def test_exact_and_incremental_agree_on_incoherent_dictionaries(ss):
    """On the BASELINE dictionaries the two residual forms differ at rounding level only: same iteration counts but for
    stop-test flips, coefficients within the gate of each other."""
    from sparsesolvers import synth
    A = synth.dictionary(256, 1024, dtype=np.float32, seed=1234 + 1)
    Y, _ = synth.sparse_signals(A, 512, 16, seed=21)
    s = ss.Homotopy(A)
    s.set_residual_mode("incremental")
    a = s.solve_batch(Y, 1e-3, 48)
    s.set_residual_mode("exact")
    b = s.solve_batch(Y, 1e-3, 48)
    c = parity.counts(parity.compare(a[0], a[1], a[2], b[0], b[1], b[2], c0_inf(A, Y)))
    print("exact vs incremental, C1 shape:", c)
    assert c["iter"] <= 8 and c["support"] == 0 and c["coef_max"] <= 1e-3


def test_max_iterations_far_beyond_the_count_ring(ss):
    """The reference takes any uint32 max_iterations.  The lockstep loop keeps its per-iteration signal counts in a ring, so a
    budget of thousands of iterations on a small dictionary (several signals: the lockstep path, not the single-signal kernel)
    runs without any per-iteration storage; an active-set bound above 4096 atoms is refused before anything is queued."""
    from oracle import port
    from sparsesolvers import synth
    A = synth.dictionary(32, 64, dtype=np.float32, seed=3)
    Y, _ = synth.sparse_signals(A, 6, 3, seed=4, normalise_sign=False)
    Xr, itr, er = port.homotopy(A, Y, 1e-3, 5000)
    s = ss.Homotopy(A)
    X, it, err = s.solve_batch(Y, 1e-3, 5000)
    assert np.isfinite(X).all() and (it >= 1).all() and (it <= 5000).all()
    short = itr <= 10
    assert (it[short] == itr[short]).all()
    X2, it2, _ = s.solve_batch(Y[short], 1e-3, 2 ** 31)          # a budget no array could be sized from
    assert (it2 == itr[short]).all()
    big = ss.Homotopy(np.eye(5000, dtype=np.float32))
    with pytest.raises(ss.SolverError, match="Invalid Argument"):
        big.solve_batch(np.zeros((2, 5000), dtype=np.float32), 1e-3, 5000)
    x, rep = big.solve(np.eye(5000, dtype=np.float32)[7], 1e-3, 4000)
    assert rep.iter == 1 and x[7] == 1


@pytest.mark.parametrize("m,n,k,max_iter", [(256, 1024, 16, 48), (512, 2048, 32, 80)])
def test_quirk_parity_set(ss, m, n, k, max_iter):
    """SURVEY.md 8(d) quirk-parity set: signals WITHOUT sign normalisation at the C1 / C2 shapes.  About half of them take the
    reference's wandering path (sign(|c|) start, homotopy-cpu.cpp:223-227) to max_iterations.  Iteration counts must match
    the reference (which signals wander, and when the others stop); coefficients are gated on the converging half only -- on the
    wandering half two CPU builds of the reference agree on every iteration count and on no more than ~55 % of the supports.
    The gate is the floor the two CPU builds show on this very set."""
    from sparsesolvers import synth
    A = synth.dictionary(m, n, dtype=np.float32, seed=1234 + m)
    Y, _ = synth.sparse_signals(A, 256, k, seed=7, normalise_sign=False)
    (Xr, itr, er), who = cpu_reference(A, Y, 1e-3, max_iter)
    X, it, err = ss.Homotopy(A).solve_batch(Y, 1e-3, max_iter)
    wander = itr == max_iter
    assert 0.3 < wander.mean() < 0.7
    conv = ~wander & (it == itr)
    floor_all = cpu_floor(A, Y, 1e-3, max_iter, (Xr, itr, er))
    assert (it != itr).sum() <= (parity.allowance(floor_all["iter"]) if floor_all else 5), f"{(it != itr).sum()} iteration-count mismatches of 256"
    floor = cpu_floor(A, Y[conv], 1e-3, max_iter, (Xr[conv], itr[conv], er[conv]))
    r = parity.compare(X[conv], it[conv], err[conv], Xr[conv], itr[conv], er[conv], c0_inf(A, Y)[conv])
    c = parity.assert_parity(r, np.float32, f"quirk set {m}x{n}, converging half [vs {who}]", noise_rate=0.15 if m == 512 else 0.02, floor=floor)
    print("PARITY quirk", m, n, c, "floor(port vs ref):", floor, "wandering:", int(wander.sum()))
    assert np.isfinite(X).all()
