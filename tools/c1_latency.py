"""C1: single-signal solves, one call at a time (256x1024 float32, k=16, tol 1e-3, max_iter 48)."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "sparse-solvers_b200"))
import sparsesolvers as ss
from sparsesolvers import synth
A = synth.dictionary(256, 1024, dtype=np.float32, seed=1235)
Y, _ = synth.sparse_signals(A, 256, 16, seed=7)
s = ss.Homotopy(A)
for b in range(8): s.solve(Y[b], 1e-3, 48)
t0 = time.perf_counter(); its = []
for b in range(256):
    x, rep = s.solve(Y[b], 1e-3, 48); its.append(rep.iter)
dt = (time.perf_counter() - t0) / 256
print(f"GPU single-signal: {dt * 1e3:.3f} ms/signal, {1 / dt:.0f} signals/s, mean iters {np.mean(its):.2f}, per iteration {dt / (np.mean(its) + 1) * 1e6:.1f} us")
try:
    from oracle import ref
    if ref.available():
        ref.set_blas_threads(1)
        t0 = time.perf_counter(); X, it, err = ref.homotopy(A, Y, 1e-3, 48, threads=1); dtc = (time.perf_counter() - t0) / 256
        print(f"CPU reference, 1 thread: {dtc * 1e3:.3f} ms/signal, iters match {(it == np.array(its)).mean():.3f}")
except Exception as e:
    print("no reference:", e)
