"""IRLS timing probe: tall float64 dictionaries (BASELINE config 4 read as M >= N)."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "sparse-solvers_b200"))
import sparsesolvers as ss
for (m, n, k, batch) in ((512, 128, 8, 2048), (2048, 512, 16, 296), (4096, 1024, 64, 148)):
    rng = np.random.default_rng(m)
    A = rng.standard_normal((m, n)) / np.sqrt(m)
    X0 = np.zeros((batch, n))
    for b in range(batch):
        X0[b, rng.choice(n, k, replace=False)] = 1 + np.abs(rng.standard_normal(k))
    Y = X0 @ A.T
    t0 = time.perf_counter(); s = ss.Irls(A); t_create = time.perf_counter() - t0
    t0 = time.perf_counter(); X, it, err, spd = s.solve_batch(Y, 1e-3, 50); t = time.perf_counter() - t0
    print(f"IRLS f64 {m}x{n} k={k} batch {batch}: create {t_create:.2f} s, solve {t:.3f} s = {batch / t:.1f} signals/s, "
          f"mean iters {it.mean():.2f}, spd_failure {spd.mean():.2f}", flush=True)
