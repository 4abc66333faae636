#!/usr/bin/env python
"""Driver of tests/studies/residual_modes.c: iteration counts of the residual-correlation alternatives on the reference's
own benchmark workload (bench.py grid_workload = homotopy_bench.cpp:14-48), against the reference form (mode 0), and, when
the reference build is present (oracle/_ref), against the unmodified reference itself.

    python tests/studies/residual_modes.py [--cells 64x256,256x1024,...] [--solves 32] [--out profiles/r2_residual_modes.json]
"""
import argparse, ctypes as C, json, os, subprocess, sys, time
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
HERE = os.path.dirname(os.path.abspath(__file__))
MODES = {0: "reference form", 1: "incremental c", 2: "incremental r", 3: "active-set r", 4: "stacked (fresh c', one-step signs)",
         5: "Gram form", 6: "Gram form < 16, stacked from 16", 7: "stacked, reciprocal-ordered candidates", 8: "reference form, FUSED x update"}


def lib():
    so = os.path.join(HERE, "_build", "libresidual_modes.so")
    src = os.path.join(HERE, "residual_modes.c")
    if not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        os.makedirs(os.path.dirname(so), exist_ok=True)
        subprocess.check_call(["gcc", "-O2", "-fopenmp", "-fPIC", "-shared", "-o", so, src, "-lm"])
    return C.CDLL(so)


def grid_workload(M, N, seed=0):
    rng = np.random.default_rng(seed)
    hay = (0.5 + 0.1 * rng.standard_normal((M, N))).astype(np.float32)
    sig = (0.5 + 0.1 * rng.standard_normal(M)).astype(np.float32)
    sig[::2] += 1.0
    return hay, sig


def run_cell(L, mode, hay, sig, solves, tol, max_iter, switch_iter=16):
    M, N = hay.shape
    it = np.zeros(solves, dtype=np.uint32)
    er = np.zeros(solves, dtype=np.float64)
    L.grid_cell(C.c_int(mode), C.c_int(switch_iter), hay.ctypes.data_as(C.c_void_p), sig.ctypes.data_as(C.c_void_p), C.c_size_t(M),
                C.c_size_t(N), C.c_size_t(solves), C.c_float(tol), C.c_uint32(max_iter), it.ctypes.data_as(C.c_void_p),
                er.ctypes.data_as(C.c_void_p))
    return it, er


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cells", default="64x256,64x1024,256x256,256x1024,512x1024,256x2048")
    ap.add_argument("--solves", type=int, default=32)
    ap.add_argument("--modes", default="0,1,2,3,4,5,6")
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    L = lib()
    from oracle import ref
    have_ref = ref.available()
    if have_ref:
        ref.set_blas_threads(1)
    out = {"what": __doc__.strip().split("\n")[0], "modes": MODES, "cells": []}
    for cell in args.cells.split(","):
        M, N = (int(v) for v in cell.split("x"))
        hay, sig = grid_workload(M, N)
        rec = {"m": M, "n": N, "solves": args.solves, "modes": {}}
        base = None
        if have_ref:
            itr = np.zeros(args.solves, dtype=np.int64)
            for i in range(args.solves):
                n = i % N
                hay[::2, n] += 1.0
                _, it1, _ = ref.homotopy(hay, sig[None, :], 0.1, N, threads=1)
                itr[i] = it1[0]
                hay[::2, n] -= 1.0
            rec["reference_build"] = {"mean_iters": float(itr.mean()), "hit_max_iter": int((itr == N).sum())}
        for mode in (int(v) for v in args.modes.split(",")):
            t0 = time.time()
            it, er = run_cell(L, mode, hay, sig, args.solves, 0.1, N)
            if mode == 0:
                base = it
            r = {"name": MODES[mode], "mean_iters": float(it.mean()), "hit_max_iter": int((it == N).sum()),
                 "iters_equal_mode0": int((it == base).sum()) if base is not None else None,
                 "within_2_of_mode0": int((np.abs(it.astype(np.int64) - base.astype(np.int64)) <= 2).sum()) if base is not None else None,
                 "seconds": round(time.time() - t0, 1)}
            if have_ref:
                r["iters_equal_reference_build"] = int((it == itr).sum())
            rec["modes"][str(mode)] = r
            print(f"{M}x{N} mode {mode} {MODES[mode]:38s} mean {r['mean_iters']:8.1f} max-iter {r['hit_max_iter']:3d} "
                  f"== mode0 {r['iters_equal_mode0']} (+-2: {r['within_2_of_mode0']})"
                  + (f" == ref {r['iters_equal_reference_build']} (ref mean {rec['reference_build']['mean_iters']:.1f}, max-iter {rec['reference_build']['hit_max_iter']})" if have_ref else ""),
                  flush=True)
        out["cells"].append(rec)
    if args.out:
        with open(args.out, "w") as f:
            json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()
