"""Timing experiments on the tcgen05 contraction: which part of the kernel bounds it (SSB200_TC_DBG bits)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "sparse-solvers_b200"))
import sparsesolvers as ss
M, N, rows = 512, 2048, 65536
rng = np.random.default_rng(0)
A = (rng.standard_normal((M, N)) / np.sqrt(M)).astype(np.float32)
P = rng.standard_normal((rows, M)).astype(np.float32)
for path in ("tcgen05", "tcgen05-tf32"):
    for dbg in (0, 1, 2, 3, 4, 5, 7):
        os.environ["SSB200_TC_DBG"] = str(dbg)
        s = ss.Homotopy(A)
        _, ms = s.debug_contract(P, path, reps=10)
        print(path, "dbg", dbg, "ms %.3f" % ms, "TF/s %.0f" % (2.0 * M * N * rows / (ms * 1e-3) / 1e12), flush=True)
        del s
