"""Golden vectors at the BASELINE config 4 shape (IRLS, float64, 4096 x 1024 -- the tall reading) from the
UNMODIFIED reference (oracle/_ref).  Minutes per signal on the CPU (the reference re-forms Q and R inside every solve,
src/solvers/irls-cpu.cpp:72-73), so this runs offline in the build container:

    python tests/golden/make_golden_c4.py [signals]

The dictionary (32 MB) is not stored: the fixture keeps the seed, checksums of A and Y, and the reference's outputs; the test
regenerates the inputs from the seed (numpy Generator, PCG64) and checks the checksums before comparing.
"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import ref                       # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))
M, N, K, SEED = 4096, 1024, 64, 4404


def inputs(batch, seed=SEED, m=M, n=N, k=K):
    """the C4 workload of bench.py (make_workload, solver == 'irls')"""
    rng = np.random.default_rng(seed)
    A = (rng.standard_normal((m, n)) / np.sqrt(m)).astype(np.float64)
    X0 = np.zeros((batch, n), dtype=np.float64)
    for b in range(batch):
        X0[b, rng.choice(n, k, replace=False)] = 1 + np.abs(rng.standard_normal(k))
    return A, (X0 @ A.T).astype(np.float64)


if __name__ == "__main__":
    batch = int(sys.argv[1]) if len(sys.argv) > 1 else 3
    A, Y = inputs(batch)
    ref.set_blas_threads(1)
    t0 = time.time()
    X, it, err, spd = ref.irls(A, Y, 1e-3, 50, threads=min(batch, len(os.sched_getaffinity(0))))
    print("reference IRLS 4096x1024:", batch, "signals in", round(time.time() - t0, 1), "s; iters", it, "spd", spd.astype(int), "eps", err)
    np.savez_compressed(os.path.join(OUT, "irls_f64_4096x1024_k64_c4.npz"), seed=np.int64(SEED), batch=np.int64(batch), m=np.int64(M),
                        n=np.int64(N), k=np.int64(K), a_sum=np.float64(A.sum()), a_abs=np.float64(np.abs(A).sum()), y_sum=np.float64(Y.sum()),
                        X=X, iters=it, errs=err, spd=spd, tol=np.float64(1e-3), max_iter=np.int64(50), blas=np.bytes_(ref.blas_config()))
