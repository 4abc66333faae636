import os, sys, numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "sparse-solvers_b200"))
import sparsesolvers as ss
m, n, k, batch = 2048, 512, 16, 296
rng = np.random.default_rng(m)
A = rng.standard_normal((m, n)) / np.sqrt(m)
X0 = np.zeros((batch, n))
for b in range(batch):
    X0[b, rng.choice(n, k, replace=False)] = 1 + np.abs(rng.standard_normal(k))
s = ss.Irls(A); X, it, err, spd = s.solve_batch(X0 @ A.T, 1e-3, 50); print(it.mean())
