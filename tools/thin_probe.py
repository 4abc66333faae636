import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "sparse-solvers_b200"))
import sparsesolvers as ss
rng = np.random.default_rng(0)
A = (rng.standard_normal((512, 2048)) / np.sqrt(512)).astype(np.float32)
s = ss.Homotopy(A)
for rows in (1, 8, 32, 33, 64, 256, 1024):
    P = rng.standard_normal((rows, 512)).astype(np.float32)
    for path in ("tcgen05", "simt"):
        _, ms = s.debug_contract(P, path, reps=50)
        print(rows, path, "us/launch %.1f" % (ms * 1e3), flush=True)
