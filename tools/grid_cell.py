"""One cell of the reference bench grid (bench.py --config GRID), a few solves, refresh and solve timed apart:
    python tools/grid_cell.py M N [solves [max_iter]]"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "sparse-solvers_b200"))
import numpy as np, bench
import sparsesolvers as ss
M, N = int(sys.argv[1]), int(sys.argv[2]); solves = int(sys.argv[3]) if len(sys.argv) > 3 else 3
max_iter = int(sys.argv[4]) if len(sys.argv) > 4 else N
hay, sig = bench.grid_workload(M, N)
s = ss.Homotopy(hay)
s.set_timing(True)
tr = ts = 0.0; its = 0
for i in range(solves + 2):
    n = i % N
    hay[::2, n] += 1.0
    t0 = time.perf_counter(); s.refresh(hay); t1 = time.perf_counter()
    x, r = s.solve(sig, 0.1, max_iter); t2 = time.perf_counter()
    hay[::2, n] -= 1.0
    if i >= 2: tr += t1 - t0; ts += t2 - t1; its += r.iter
    if i == 1: s.reset_counters()
c = s.counters(); k = s.timing()
print(f"{M}x{N}: refresh {tr / solves * 1e6:.0f} us, solve {ts / solves * 1e6:.0f} us wall ({k['step_ms'] / solves * 1e3:.0f} us in step kernels, {k['contraction_ms'] / solves * 1e3:.0f} us contraction), "
      f"{its / solves:.1f} iterations, {ts / max(its, 1) * 1e6:.1f} us per iteration, launches per solve {c['launches'] / solves:.1f}, residual {s.residual_info()}")
