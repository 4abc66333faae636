"""One cell of the reference bench grid (bench.py --config GRID), a few solves:  python tools/grid_cell.py M N [solves]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "sparse-solvers_b200"))
import numpy as np, bench
import sparsesolvers as ss
M, N = int(sys.argv[1]), int(sys.argv[2]); solves = int(sys.argv[3]) if len(sys.argv) > 3 else 3
max_iter = int(sys.argv[4]) if len(sys.argv) > 4 else N
hay, sig = bench.grid_workload(M, N)
s = ss.Homotopy(hay)
for i in range(solves):
    n = i % N
    hay[::2, n] += 1.0
    s.refresh(hay)
    x, r = s.solve(sig, 0.1, max_iter)
    hay[::2, n] -= 1.0
    print(M, N, i, r, int(np.argmax(np.abs(x))), flush=True)
