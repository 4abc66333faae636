#!/usr/bin/env python
"""Probe (GPU box): the reference bench workload in FLOAT64, GPU against the float64 restatement.  In double precision the
path is far less chaotic, so iteration counts and coefficients should agree on most solves; a kernel bug in the large-active-set
variants (K up to M, inverse worked on in HBM) would show as early divergence.  CPU sides are cached (tests/_parity_cache).
    python tests/studies/grid_f64_probe.py --cpu | --gpu
"""
import argparse, os, sys, json
import numpy as np
HERE = os.path.dirname(os.path.abspath(__file__)); ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "sparse-solvers_b200"))
CACHE = os.path.join(ROOT, "tests", "_parity_cache")
CELLS = [(64, 256, 8), (256, 256, 8), (64, 2048, 8), (256, 1024, 8), (512, 1024, 4)]


def workload(M, N):
    rng = np.random.default_rng(0)
    hay = (0.5 + 0.1 * rng.standard_normal((M, N)))
    sig = (0.5 + 0.1 * rng.standard_normal(M))
    sig[::2] += 1.0
    return hay, sig


ap = argparse.ArgumentParser(); ap.add_argument("--cpu", action="store_true"); ap.add_argument("--gpu", action="store_true"); ap.add_argument("--max-iter", type=int, default=0)
args = ap.parse_args()
for M, N, solves in CELLS:
    hay, sig = workload(M, N)
    mi = args.max_iter or N
    path = os.path.join(CACHE, f"grid_f64_{M}x{N}_{mi}.npz")
    if args.cpu:
        from oracle import port
        X = np.zeros((solves, N)); it = np.zeros(solves, dtype=np.int64); er = np.zeros(solves)
        for i in range(solves):
            hay[::2, i] += 1.0
            x, t, e = port.homotopy(hay, sig[None, :], 0.1, mi)
            hay[::2, i] -= 1.0
            X[i], it[i], er[i] = x[0], t[0], e[0]
        os.makedirs(CACHE, exist_ok=True)
        np.savez_compressed(path, X=X, it=it, er=er)
        print(M, N, "port f64 iters", it.tolist())
    if args.gpu:
        import sparsesolvers as ss
        z = np.load(path)
        s = ss.Homotopy(hay)
        for mode in ("exact", "incremental"):
            s.set_residual_mode(mode)
            its, devs = [], []
            for i in range(solves):
                hay[::2, i] += 1.0
                s.refresh(hay)
                x, rep = s.solve(sig, 0.1, mi)
                hay[::2, i] -= 1.0
                its.append(rep.iter)
                devs.append(float(np.abs(x - z["X"][i]).max() / max(1e-300, np.abs(z["X"][i]).max())))
            print(f"{M}x{N} max_iter {mi} [{mode}] gpu iters {its} port {z['it'].tolist()} max rel dev {['%.1e' % d for d in devs]}", flush=True)
