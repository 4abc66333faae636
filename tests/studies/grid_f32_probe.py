#!/usr/bin/env python
"""Probe (GPU box): float32 reference-bench cells, GPU against the reference build, for a list of max_iterations (the active-set
capacity min(max_iterations + 1, N) selects the step-kernel variant) and residual modes.
    python tests/studies/grid_f32_probe.py 256x1024 --max-iters 1024,700 --solves 32
"""
import argparse, os, sys
import numpy as np
HERE = os.path.dirname(os.path.abspath(__file__)); ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "sparse-solvers_b200"))
import bench
from oracle import ref
import sparsesolvers as ss

ap = argparse.ArgumentParser(); ap.add_argument("cell"); ap.add_argument("--max-iters", default="0"); ap.add_argument("--solves", type=int, default=32)
ap.add_argument("--modes", default="exact")
args = ap.parse_args()
M, N = (int(v) for v in args.cell.split("x"))
ref.set_blas_threads(1)
for mi in (int(v) or N for v in args.max_iters.split(",")):
    hay, sig = bench.grid_workload(M, N)
    s = ss.Homotopy(hay)
    itr = []
    for i in range(args.solves):
        hay[::2, i % N] += 1.0
        itr.append(int(ref.homotopy(hay, sig[None, :], 0.1, mi, threads=1)[1][0]))
        hay[::2, i % N] -= 1.0
    itr = np.array(itr)
    for mode in args.modes.split(","):
        s.set_residual_mode(mode)
        itg = []
        for i in range(args.solves):
            hay[::2, i % N] += 1.0
            s.refresh(hay)
            itg.append(s.solve(sig, 0.1, mi)[1].iter)
            hay[::2, i % N] -= 1.0
        itg = np.array(itg)
        print(f"{M}x{N} max_iter {mi} [{mode}] hit-max gpu {(itg == mi).sum()} ref {(itr == mi).sum()}  mean gpu {itg.mean():.1f} ref {itr.mean():.1f}  equal {(itg == itr).sum()}/{args.solves}", flush=True)
        print("   gpu", itg.tolist()); print("   ref", itr.tolist(), flush=True)
