#!/usr/bin/env python
"""Parity evidence for the float32 homotopy path, reproducible: mismatch counts of every pair side by side.

    python tests/noise_floor.py --cpu                  # here (no GPU): reference build vs its plain-C restatement
    python tests/noise_floor.py --gpu [--tag NAME]     # on the GPU box: this library vs the reference build / the restatement
    python tests/noise_floor.py --merge OUT.json       # fold tests/_parity_cache/*.json into one record (profiles/r2_parity.json)

For every configuration the same seeded inputs go to
  ref   the UNMODIFIED reference (oracle/_ref, OpenBLAS through the reference's own dlopen)
  port  the plain-C restatement (oracle/ss_oracle.c)
  gpu   libssb200.so through the C ABI, in the residual modes asked for (auto / exact / incremental: the batched lockstep path;
        single / single-exact: one call per signal, i.e. the single-signal cluster kernel)
and the pairs (port, ref) = CPU-against-CPU noise floor, (gpu, ref), (gpu, port) are compared with tests/parity.py:
iter = iteration-count mismatches, support = hard support mismatches, soft = support flips of rounding residue,
coef_1e-4 / coef_1e-3 = signals whose largest coefficient error relative to max|x_ref| exceeds that, err = solution_error
mismatches.  Counts other than `iter` are over the signals whose iteration counts agree.

The CPU results are cached under tests/_parity_cache/ (git-ignored, travels to the GPU box), so the GPU box spends its time
on the GPU side only.  Lives under tests/ because it drives the oracle (only tests/, smoke() and bench.py's CPU legs may).
"""
import argparse
import glob
import json
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "sparse-solvers_b200"))
sys.path.insert(0, HERE)
CACHE = os.path.join(HERE, "_parity_cache")

import parity  # noqa: E402


def configs(quick=False):
    from sparsesolvers import synth
    n = 256 if quick else 1024

    def gauss(m, nn, k, mi, count, seed, norm=True):
        def make():
            A = synth.dictionary(m, nn, dtype=np.float32, seed=1234 + m)
            Y, _ = synth.sparse_signals(A, count, k, seed=seed, normalise_sign=norm)
            return A, Y
        return dict(make=make, tol=1e-3, max_iter=mi, n=count)

    def face(mi):
        def make():
            A, Y, _ = synth.face_recognition(batch=n, seed=77)
            return A, Y
        return dict(make=make, tol=1e-3, max_iter=mi, n=n)
    return {
        "C1 256x1024 k=16": gauss(256, 1024, 16, 48, n, 11),
        "C2 512x2048 k=32": gauss(512, 2048, 32, 80, n, 12),
        "C3 504x1216 face @60": face(60),
        "C3 504x1216 face @100": face(100),
        "C5 4096x16384 k=128": gauss(4096, 16384, 128, 272, 16 if quick else 64, 15),
        "quirk C1 (no sign normalisation)": gauss(256, 1024, 16, 48, n, 7, norm=False),
        "quirk C2 (no sign normalisation)": gauss(512, 2048, 32, 80, n, 7, norm=False),
    }


def counts(r):
    good = ~r["iter"]
    rel = r["coef_rel"]
    return {"n": int(len(good)), "iter": int(r["iter"].sum()), "support": int((r["support"] & good).sum()),
            "soft": int((r["support_soft"] & good).sum()), "coef_1e-4": int(((rel > 1e-4) & good).sum()),
            "coef_1e-3": int(((rel > 1e-3) & good).sum()), "err": int((r["err"] & good).sum()),
            "max_coef_rel": float(rel[good].max()) if good.any() else None}


def compare(a, b, c0, max_iter=None):
    """counts of a against b; with max_iter (the quirk sets): iteration counts over all signals, everything else over the
    signals b CONVERGES on -- on the half that wanders to max_iter (sign(|c|) start, homotopy-cpu.cpp:223-227) the path is
    chaotic and two CPU builds of the reference agree on the iteration count only"""
    if max_iter is None:
        return counts(parity.compare(a[0], a[1], a[2], b[0], b[1], b[2], c0))
    conv = np.asarray(b[1]) < max_iter
    out = counts(parity.compare(a[0][conv], a[1][conv], a[2][conv], b[0][conv], b[1][conv], b[2][conv], c0[conv]))
    out["iter_all"] = int((np.asarray(a[1]) != np.asarray(b[1])).sum())
    out["n_all"] = int(len(conv))
    out["wandering"] = int((~conv).sum())
    return out


def cache_path(name, who):
    return os.path.join(CACHE, name.replace(" ", "_").replace("/", "_") + f".{who}.npz")


def cpu_side(name, cfg, who, threads):
    """(X, iters, errs) of the reference build / the restatement, cached."""
    p = cache_path(name, who)
    if os.path.exists(p):
        z = np.load(p)
        if int(z["n"]) == cfg["n"] and int(z["max_iter"]) == cfg["max_iter"]:
            return (z["X"], z["it"], z["err"]), float(z["seconds"])
    from oracle import port, ref
    A, Y = cfg["make"]()
    t0 = time.time()
    if who == "ref":
        ref.set_blas_threads(1)
        X, it, err = ref.homotopy(A, Y, cfg["tol"], cfg["max_iter"], threads=threads)
    else:
        X, it, err = port.homotopy(A, Y, cfg["tol"], cfg["max_iter"])
    dt = time.time() - t0
    os.makedirs(CACHE, exist_ok=True)
    np.savez_compressed(p, X=X, it=it, err=err, n=cfg["n"], max_iter=cfg["max_iter"], seconds=dt)
    return (X, it, err), dt


def c0_inf(A, Y):
    return np.abs(np.asarray(Y, dtype=np.float64) @ np.asarray(A, dtype=np.float64)).max(axis=1)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cpu", action="store_true")
    ap.add_argument("--gpu", action="store_true")
    ap.add_argument("--merge", default=None)
    ap.add_argument("--tag", default="default", help="names this library build in the record (e.g. ieee-order)")
    ap.add_argument("--modes", default="auto,exact,incremental,single,single-exact",
                    help="residual modes of the batched path; single / single-exact = one ssb200_homotopy_solve call per signal")
    ap.add_argument("--only", default=None, help="substring filter on the configuration names")
    ap.add_argument("--quick", action="store_true")
    args = ap.parse_args()
    threads = len(os.sched_getaffinity(0))
    cfgs = {k: v for k, v in configs(args.quick).items() if not args.only or args.only in k}

    if args.cpu:
        from oracle import ref
        rec = {"what": "CPU against CPU: plain-C restatement vs the unmodified reference build, same inputs", "blas": ref.blas_config(), "configs": {}}
        for name, cfg in cfgs.items():
            A, Y = cfg["make"]()
            r, tr = cpu_side(name, cfg, "ref", threads)
            p, tp = cpu_side(name, cfg, "port", threads)
            rec["configs"][name] = {"port_vs_ref": compare(p, r, c0_inf(A, Y), cfg["max_iter"] if name.startswith("quirk") else None), "ref_seconds": round(tr, 1), "port_seconds": round(tp, 1),
                                    "ref_mean_iters": float(r[1].mean()), "ref_hit_max_iter": int((r[1] == cfg["max_iter"]).sum())}
            print(name, rec["configs"][name], flush=True)
        os.makedirs(CACHE, exist_ok=True)
        json.dump(rec, open(os.path.join(CACHE, "cpu.json"), "w"), indent=1)

    if args.gpu:
        import sparsesolvers as ss
        rec = {"what": "libssb200.so against the CPU results, per residual mode", "tag": args.tag,
               "lib": os.environ.get("SSB200_LIB", "sparse-solvers_b200/csrc/libssb200.so"), "configs": {}}
        for name, cfg in cfgs.items():
            A, Y = cfg["make"]()
            c0 = c0_inf(A, Y)
            r, _ = cpu_side(name, cfg, "ref", threads)
            p, _ = cpu_side(name, cfg, "port", threads)
            solver = ss.Homotopy(A)
            out = {}
            for mode in args.modes.split(","):
                single = mode.startswith("single")
                solver.set_residual_mode({"single": "auto", "single-exact": "exact"}.get(mode, mode))
                eff, mu = solver.residual_info()
                t0 = time.time()
                if single:      # one call per signal: the single-signal kernel (thread-block cluster form)
                    if A.shape[0] * A.shape[1] > (1 << 21):
                        continue
                    X = np.empty((len(Y), A.shape[1]), dtype=A.dtype); it = np.empty(len(Y), dtype=np.uint32); er = np.empty(len(Y))
                    for b in range(len(Y)):
                        X[b], rep = solver.solve(Y[b], cfg["tol"], cfg["max_iter"])
                        it[b], er[b] = rep.iter, rep.solution_error
                    g = (X, it, er)
                else:
                    g = solver.solve_batch(Y, cfg["tol"], cfg["max_iter"])
                mi_q = cfg["max_iter"] if name.startswith("quirk") else None
                out[mode] = {"mode_in_effect": eff, "coherence": round(mu, 4), "gpu_vs_ref": compare(g, r, c0, mi_q), "gpu_vs_port": compare(g, p, c0, mi_q),
                             "seconds": round(time.time() - t0, 2), "gpu_mean_iters": float(g[1].mean()),
                             "gpu_hit_max_iter": int((g[1] == cfg["max_iter"]).sum())}
                print(name, mode, out[mode], flush=True)
            out["port_vs_ref"] = compare(p, r, c0, cfg["max_iter"] if name.startswith("quirk") else None)
            rec["configs"][name] = out
            del solver
        os.makedirs(CACHE, exist_ok=True)
        json.dump(rec, open(os.path.join(CACHE, f"gpu_{args.tag}.json"), "w"), indent=1)
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        json.dump(rec, open(os.path.join(ROOT, "gpurun_out", f"parity_gpu_{args.tag}.json"), "w"), indent=1)

    if args.merge:
        merged = {"what": __doc__.strip().split("\n")[0], "records": {}}
        for src in [os.path.join(CACHE, "*.json"), os.path.join(ROOT, "gpurun_out", "parity_gpu_*.json")]:
            for f in sorted(glob.glob(src)):
                merged["records"][os.path.basename(f)[:-5]] = json.load(open(f))
        json.dump(merged, open(args.merge, "w"), indent=1)
        print("wrote", args.merge)


if __name__ == "__main__":
    main()
