"""BASELINE config 1 through the reference's OWN pybind11 module (binding.cpp, unchanged) built on libssb200 (dropin/Makefile):
one solve() per signal, wall clock; prints one JSON line.  bench.py --config C1 runs this as a subprocess (the module shares its
package name with the ctypes mirror, so the two cannot live in one process) and reports it as `dropin`.
    python dropin/time_c1.py M N k tol max_iter batch steps"""
import json
import os
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, os.path.join(HERE, "_build"))
import numpy as np           # noqa: E402
import sparsesolvers as ss   # noqa: E402  (the pybind11 module)

assert ss.__file__.startswith(os.path.join(HERE, "_build")), ss.__file__
import importlib.util          # noqa: E402
spec = importlib.util.spec_from_file_location("ssb200_synth", os.path.join(ROOT, "sparse-solvers_b200", "sparsesolvers", "synth.py"))
synth = importlib.util.module_from_spec(spec)      # the workload generator of bench.py (make_workload), loaded as a plain file
spec.loader.exec_module(synth)

M, N, k, tol, max_iter, batch, steps = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), float(sys.argv[4]), int(sys.argv[5]), int(sys.argv[6]), int(sys.argv[7])
A = synth.dictionary(M, N, dtype=np.float32, seed=1234 + 2)
Y, _ = synth.sparse_signals(A, batch, k, seed=4321)
solver = ss.Homotopy(A)
for b in range(min(batch, 48)):
    solver.solve(Y[b], tol, max_iter)
its = np.empty(batch)
t0 = time.perf_counter()
for _ in range(steps):
    for b in range(batch):
        x, rep = solver.solve(Y[b], tol, max_iter)
        its[b] = rep.iter
secs = time.perf_counter() - t0
print(json.dumps({"value": batch * steps / secs, "unit": "signals/s", "us_per_solve": 1e6 * secs / (batch * steps), "mean_iters": float(its.mean())}))
