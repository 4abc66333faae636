#!/usr/bin/env python
"""bench.py -- homotopy signals/sec on B200 (BASELINE.json metric), one JSON line on stdout.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--config C2]

A step = one pass of the hot path over one batch.  C2 as BASELINE.json states it: 65,536 k=32 signals against a
512x2048 float32 dictionary, SHARDED over the N GPUs (strong scaling: total work fixed as N grows).  The dictionary
is broadcast once over NCCL, every rank solves its contiguous slice, and inside the timed region the coefficients --
in their sparse (index, value, count) form -- and the reports are gathered to rank 0 over NCCL; nothing else is
exchanged.  --scaling weak keeps `batch` signals per GPU instead (the figure is also added to the strong line as `weak`).

  value     signals/s, inputs resident in HBM (device-pointer entry of the C ABI), CUDA-event timed
  e2e       the same through the host-pointer entry: pinned host Y in, host X + reports out, every step
  roofline  the contraction kernel Q = A^T P: algorithmic flops 2*M*N per signal-row / its measured time
  cpu_baseline  the unmodified reference (oracle/_ref) on this box's host cores, bounded sample

--impl reference times the reference's own CPU implementation (oracle/_ref, else the oracle port).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "sparse-solvers_b200"))

import numpy as np  # noqa: E402

CONFIGS = {
    # BASELINE.json configs[]; C2 is the configuration the metric is quoted on
    "C1": dict(m=256, n=1024, k=16, dtype="float32", batch=1024, tol=1e-3, max_iter=48, solver="single"),
    "C2": dict(m=512, n=2048, k=32, dtype="float32", batch=65536, tol=1e-3, max_iter=80),
    "C3": dict(m=504, n=1216, k=32, dtype="float32", batch=262144, tol=1e-3, max_iter=100, synth="face"),
    # BASELINE says 1024x4096; the reference's IRLS is undefined for M < N: the tall reading is the oracle-able one
    "C4": dict(m=4096, n=1024, k=64, dtype="float64", batch=16384, tol=1e-3, max_iter=50, solver="irls"),
    "C5": dict(m=4096, n=16384, k=128, dtype="float32", batch=8192, tol=1e-3, max_iter=272),
    # the reference's own benchmark grid (src/solvers/homotopy_bench.cpp:57-60), SURVEY.md §8(f) N2
    "GRID": dict(m=512, n=2048, k=0, dtype="float32", batch=32, tol=0.1, max_iter=2048, solver="grid"),
}


# stdout carries the ONE JSON line and nothing else: whatever libraries print there (NCCL's version banner is a plain printf)
# is sent to stderr by pointing fd 1 at fd 2 for the whole run; the line itself goes to the saved descriptor.
_REAL_STDOUT = os.fdopen(os.dup(1), "w")
os.dup2(2, 1)


def emit(line):
    _REAL_STDOUT.write(json.dumps(line) + "\n")
    _REAL_STDOUT.flush()


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=None, help="timed steps (default 20; 3 for the IRLS config and the reference arm)")
    ap.add_argument("--warmup", type=int, default=None, help="untimed warm-up steps (default 5; 3 for IRLS / reference)")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="C2", choices=sorted(CONFIGS))
    ap.add_argument("--batch", type=int, default=None, help="signals per GPU per step (default: the config's)")
    ap.add_argument("--path", default="auto", choices=["auto", "simt", "tcgen05", "tcgen05-tf32"])
    ap.add_argument("--cpu-sample", type=int, default=None, help="signals in the cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cells", default=None, help="GRID: only these cells, e.g. 64x256,256x1024")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-kernel-timing", action="store_true", help="no per-kernel CUDA events in the timed region (roofline absent)")
    ap.add_argument("--gather-x", action="store_true", help="gather the DENSE coefficients instead of the sparse form")
    ap.add_argument("--scaling", default="strong", choices=["strong", "weak"], help="N > 1: shard the config's batch (strong, default) or keep it per GPU (weak)")
    ap.add_argument("--no-weak", action="store_true", help="N > 1, strong: do not add the weak-scaling figure")
    ap.add_argument("--residual", default="auto", choices=["auto", "exact", "incremental"], help="residual-correlation mode of the handle")
    ap.add_argument("--no-exact-extra", action="store_true", help="do not time the other residual mode as an extra key")
    args = ap.parse_args()
    slow = args.impl == "reference" or CONFIGS[args.config].get("solver") == "irls"
    if args.steps is None:
        args.steps = 3 if slow else 20
    if args.warmup is None:
        args.warmup = 3 if slow else 5
    return args


def make_workload(cfg, batch, seed):
    from sparsesolvers import synth
    dt = np.dtype(cfg["dtype"])
    if cfg.get("synth") == "face":
        A, Y, _ = synth.face_recognition(cfg["m"], cfg["n"], batch=batch, seed=seed, dtype=dt)
        return A, Y
    if cfg.get("solver") == "irls":
        rng = np.random.default_rng(seed)
        A = (rng.standard_normal((cfg["m"], cfg["n"])) / np.sqrt(cfg["m"])).astype(dt)
        X0 = np.zeros((batch, cfg["n"]), dtype=dt)
        for b in range(batch):
            X0[b, rng.choice(cfg["n"], cfg["k"], replace=False)] = 1 + np.abs(rng.standard_normal(cfg["k"]))
        return A, (X0 @ A.T).astype(dt)
    A = synth.dictionary(cfg["m"], cfg["n"], dtype=dt, seed=1234 + 2)
    Y, X0 = synth.sparse_signals(A, batch, cfg["k"], seed=seed)
    return A, Y


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def cpu_reference_solve(A, Y, cfg, threads, full=False):
    """The reference's CPU path on `threads` host threads (one BLAS thread each: its best throughput mode,
    BASELINE.md §3).  Returns (seconds, kind, iters) (+ X, errs with full=True)."""
    from oracle import ref, port
    if ref.available() and os.path.exists("/root/repo/oracle/_ref/libcblas_fwd.so"):
        ref.set_blas_threads(1)
        t0 = time.perf_counter()
        X, it, err = ref.homotopy(A, Y, cfg["tol"], cfg["max_iter"], threads=threads)
        secs, kind = time.perf_counter() - t0, "reference"
    else:
        os.environ.setdefault("OMP_NUM_THREADS", str(threads))
        t0 = time.perf_counter()
        X, it, err = port.homotopy(A, Y, cfg["tol"], cfg["max_iter"])
        secs, kind = time.perf_counter() - t0, "port"
    return (secs, kind, it, X, err) if full else (secs, kind, it)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", os.environ.get("BENCH_CLOCK_MS", "20")],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), [c.strip() for c in line.split(",")]))

    def window(self, t0, t1):
        """keep the samples taken inside the timed region (the sampler is started before the warm-up: the first
        nvidia-smi query on a fresh box takes seconds and stalls CUDA calls while it initialises)"""
        inside = [r for t, r in self.rows if t0 <= t <= t1]
        self.rows = inside if inside else [r for _, r in self.rows[-3:]]

    def stop(self):
        if not self.proc:
            return None
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        self.rows = [r[1] if isinstance(r, tuple) else r for r in self.rows]
        sm = [float(r[0]) for r in self.rows if len(r) >= 7 and r[0].replace(".", "").isdigit()]
        if not sm:
            return None
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) >= 7 and r[3 + i].lower().startswith("active") for r in self.rows)]
        mx = [float(r[1]) for r in self.rows if len(r) >= 7 and r[1].replace(".", "").isdigit()]
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm)}


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)), "measured"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback"


def ncu_traffic(config):
    """DRAM bytes per full-chunk launch (dram__bytes_read.sum + dram__bytes_write.sum) of each kernel, from the committed
    `ncu --set full` capture of this same command (profiles/ncu_traffic.json, written by tools/ncu_summary.py --traffic);
    captures exist for the configurations listed there only."""
    p = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if not os.path.exists(p):
        return {}
    with open(p) as f:
        return json.load(f).get(config, {}).get("bytes_per_launch", {})


def run_reference(args, cfg):
    """--impl reference: the reference's CPU solver on this box's host cores, bounded sample per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = host_threads()
    sample = args.cpu_sample or (256 if args.config == "C2" else 64)
    A, Y = make_workload(cfg, sample, seed=4321)
    for _ in range(args.warmup):
        cpu_reference_solve(A, Y[: max(8, sample // 8)], cfg, threads)
    secs, kind = [], "port"
    for _ in range(args.steps):
        s, kind, it = cpu_reference_solve(A, Y, cfg, threads)
        secs.append(s)
    total = float(sum(secs))
    value = sample * args.steps / total
    line = {
        "impl": "reference", "metric": "homotopy signals/sec", "value": value, "unit": "signals/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps,
        "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{args.config}: homotopy A {cfg['m']}x{cfg['n']} {cfg['dtype']}, {cfg['batch']} k={cfg['k']} signals "
                               + ("sharded over the GPUs" if (args.gpus > 1 and args.scaling == "strong") else "per GPU")
                               + f", tol {cfg['tol']}, max_iter {cfg['max_iter']}",
                   "batch_per_step": sample, "note": "the CPU arm times a bounded sample of the same workload per step"},
        "cpu_baseline": {"value": value, "unit": "signals/s", "cores": threads, "kind": kind,
                         "sample": f"{sample} signals of the {args.config} workload per step, {threads} host threads x 1 BLAS thread"},
        "e2e": {"value": value, "unit": "signals/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


def run_irls(args, cfg):
    """C4: IRLS float64 through the host-pointer entry (the only one the IRLS handle has): one CTA per signal for its
    whole solve, so `value` and `e2e` are the same measurement.  Single GPU."""
    import torch
    import sparsesolvers as ss
    batch = args.batch or cfg["batch"]
    A, Y = make_workload(cfg, batch, seed=4321)
    solver = ss.Irls(A)
    for _ in range(args.warmup):
        solver.solve_batch(Y[: min(batch, 592)], cfg["tol"], cfg["max_iter"])
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        X, it, err, spd = solver.solve_batch(Y, cfg["tol"], cfg["max_iter"])
    secs = time.perf_counter() - t0
    M, N = cfg["m"], cfg["n"]
    iters = float(it.mean()) + float(spd.mean())     # a failing iteration still forms and factors G
    flops = (1.0 * N * N * M + N ** 3 / 3.0) * iters * batch * args.steps     # mul+add pairs counted as 2: N^2/2*M*2 + N^3/6*2
    value = batch * args.steps / secs
    # FP64 peak measured on this device (ssb200_fp64_peak_probe): independent DFMA chains and mma.sync.m8n8k4.f64 tiles.  The solve
    # kernel rounds every multiply and add separately (it has to reproduce the reference's SPD verdicts), so it cannot use FMA: its
    # ceiling is HALF the DFMA peak, which is the denominator used here.
    dfma, dmma = ss.fp64_peak("dfma"), ss.fp64_peak("dmma")
    peak = dfma / 2.0
    cpu = None
    if not args.no_cpu_baseline:
        from oracle import ref
        if ref.available():
            threads = host_threads()
            sample = args.cpu_sample or min(8, threads)
            ref.set_blas_threads(1)
            t1 = time.perf_counter()
            Xr, itr, er, sr = ref.irls(A, Y[:sample], cfg["tol"], cfg["max_iter"], threads=min(sample, threads))
            dt = time.perf_counter() - t1
            cpu = {"value": sample / dt, "unit": "signals/s", "cores": min(sample, threads), "kind": "reference",
                   "sample": f"first {sample} signals, one per host thread x 1 BLAS thread, {dt:.1f} s (the reference re-forms Q and R inside every solve, irls-cpu.cpp:72-73)",
                   "iters_match_gpu": int((itr == it[:sample]).sum()), "spd_failure_match_gpu": int((sr == spd[:sample]).sum()),
                   "solution_error_max_rel_dev": float(np.max(np.abs(er - err[:sample]) / np.abs(er)))}
    line = {"metric": "irls signals/sec", "value": value, "unit": "signals/s", "n_gpus": 1, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * secs / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": {"workload": f"C4 (tall reading): IRLS A {M}x{N} float64, {batch} k={cfg['k']} signals, tol {cfg['tol']}, max_iter {cfg['max_iter']}",
                       "mean_iters": float(it.mean()), "spd_failure_frac": float(spd.mean())},
            "e2e": {"value": value, "unit": "signals/s", "h2d_bytes_per_step": int(Y.nbytes), "d2h_bytes_per_step": int(X.nbytes + it.nbytes + err.nbytes + spd.nbytes)},
            "gpu_launches": solver.counters()["launches"] if hasattr(solver, "counters") else None,
            "roofline": {"kernel": "irls_solve_kernel (Gram + Cholesky tiles)", "bound": "fp64", "achieved": flops / secs / 1e12, "peak": peak,
                         "unit": "TFLOP/s", "frac": flops / secs / 1e12 / peak, "traffic": None,
                         "peak_source": f"measured on this device: DFMA {dfma:.1f} TFLOP/s, DMMA {dmma:.1f} TFLOP/s (ssb200_fp64_peak_probe); separately "
                                        "rounded mul / add cannot use FMA: half the DFMA figure is the ceiling"},
            "cpu_baseline": cpu}
    emit(line)


def grid_workload(M, N, seed=0):
    """The reference bench's inputs (src/solvers/homotopy_bench.cpp:14-48): haystack ~ N(0.5, 0.1) M x N, signal ~ N(0.5, 0.1)
    with +1 on every second entry; before solve i the same +1 pattern is added to column i % N and taken away after."""
    rng = np.random.default_rng(seed)
    hay = (0.5 + 0.1 * rng.standard_normal((M, N))).astype(np.float32)
    sig = (0.5 + 0.1 * rng.standard_normal(M)).astype(np.float32)
    sig[::2] += 1.0
    return hay, sig


def run_grid(args, cfg):
    """The reference's own benchmark grid (homotopy_bench.cpp:57-60: M 16..512 x N 16..2048, multiplier 4, tolerance 0.1,
    max_iter = N), re-hosted without Google Benchmark: per cell, `solves` single-signal solves with the dictionary edited in
    place between them (the handle follows the edit through ssb200_homotopy_refresh), GPU next to the reference on one
    host thread.  One JSON line; `value` = solves per second over the whole grid."""
    import torch
    import sparsesolvers as ss
    from oracle import ref
    Ms, Ns = (16, 64, 256, 512), (16, 64, 256, 1024, 2048)
    solves = args.batch or 32
    cells, total_gpu, total_n = [], 0.0, 0
    have_ref = (not args.no_cpu_baseline) and ref.available()
    if have_ref:
        ref.set_blas_threads(1)
    only = None if not args.cells else {tuple(int(v) for v in c.split("x")) for c in args.cells.split(",")}
    for M in Ms:
        for N in Ns:
            if only is not None and (M, N) not in only:
                continue
            hay, sig = grid_workload(M, N)
            solver = ss.Homotopy(hay)
            it_gpu, it_ref = np.zeros(solves, dtype=np.int64), np.zeros(solves, dtype=np.int64)
            for rep in range(2):                       # first pass warms the handle up
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                for i in range(solves):
                    n = i % N
                    hay[::2, n] += 1.0
                    solver.refresh(hay)
                    x, r = solver.solve(sig, 0.1, N)
                    it_gpu[i] = r.iter
                    hay[::2, n] -= 1.0
                gpu_s = time.perf_counter() - t0
            cell = {"m": M, "n": N, "gpu_ms_per_solve": 1e3 * gpu_s / solves, "gpu_mean_iters": float(it_gpu.mean())}
            if have_ref:
                t0 = time.perf_counter()
                done = 0
                for i in range(solves):                 # bounded: the largest cells take ~20 s per solve on one core
                    n = i % N
                    hay[::2, n] += 1.0
                    xr, itr, er = ref.homotopy(hay, sig[None, :], 0.1, N, threads=1)
                    it_ref[i] = itr[0]
                    hay[::2, n] -= 1.0
                    done += 1
                    if time.perf_counter() - t0 > 12.0:
                        break
                cpu_s = time.perf_counter() - t0
                cell.update(cpu_ms_per_solve=1e3 * cpu_s / done, cpu_solves=done, cpu_mean_iters=float(it_ref[:done].mean()),
                            gpu_mean_iters_same_solves=float(it_gpu[:done].mean()),
                            iters_equal=int((it_ref[:done] == it_gpu[:done]).sum()),
                            gpu_hit_max_iter=int((it_gpu[:done] == N).sum()), cpu_hit_max_iter=int((it_ref[:done] == N).sum()),
                            gpu_hit_max_iter_all_solves=int((it_gpu == N).sum()))
            cells.append(cell)
            total_gpu += gpu_s
            total_n += solves
            del solver
    line = {"metric": "homotopy solves/sec over the reference bench grid", "value": total_n / total_gpu, "unit": "solves/s", "n_gpus": 1,
            "steps": 1, "warmup": 1, "ms_per_step": 1e3 * total_gpu, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"GRID: homotopy_bench.cpp grid, M {Ms} x N {Ns}, tol 0.1, max_iter N, {solves} solves per cell, "
                                   "dictionary edited in place and refreshed before every solve"},
            "cells": cells, "cpu_baseline": {"kind": "reference", "cores": 1, "unit": "ms per solve, per cell (cells[].cpu_ms_per_solve)",
                                             "sample": f"{solves} solves per cell"} if have_ref else None}
    emit(line)


def run_single(args, cfg):
    """C1: the single-signal latency path -- `batch` signals solved ONE CALL AT A TIME through the reference-shaped
    solve(y) entry (host buffers in and out, so `value` and `e2e` are the same measurement)."""
    import torch
    import sparsesolvers as ss
    batch = args.batch or cfg["batch"]
    A, Y = make_workload(cfg, batch, seed=4321)
    solver = ss.Homotopy(A)
    for b in range(min(batch, 16 * max(1, args.warmup))):
        solver.solve(Y[b], cfg["tol"], cfg["max_iter"])
    torch.cuda.synchronize()
    its = np.empty(batch)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        for b in range(batch):
            x, rep = solver.solve(Y[b], cfg["tol"], cfg["max_iter"])
            its[b] = rep.iter
    secs = time.perf_counter() - t0
    M, N = cfg["m"], cfg["n"]
    value = batch * args.steps / secs
    peaks, peak_src = measured_peaks()
    nbytes = 4.0 * M * N * (its.mean() + 1) * batch * args.steps            # one pass over A per iteration + the initial one
    cpu = None
    if not args.no_cpu_baseline:
        sample = min(batch, args.cpu_sample or 256)
        from oracle import ref
        if ref.available():
            ref.set_blas_threads(1)
            t1 = time.perf_counter(); Xr, itr, er = ref.homotopy(A, Y[:sample], cfg["tol"], cfg["max_iter"], threads=1); dt = time.perf_counter() - t1
            cpu = {"value": sample / dt, "unit": "signals/s", "cores": 1, "kind": "reference",
                   "sample": f"first {sample} signals, one call at a time, 1 thread, {dt:.2f} s", "iters_match_gpu": int((itr == its[:sample]).sum())}
    # the same calls through the reference's own pybind11 module linked on libssb200 (dropin/): what a user of the reference sees after
    # swapping the library -- includes the stub's per-solve check that the dictionary view has not been edited
    dropin = None
    mod = os.path.join(ROOT, "dropin", "_build", "sparsesolvers")
    if os.path.isdir(mod):
        import subprocess
        r = subprocess.run([sys.executable, os.path.join(ROOT, "dropin", "time_c1.py"), str(cfg["m"]), str(cfg["n"]), str(cfg["k"]), str(cfg["tol"]),
                            str(cfg["max_iter"]), str(batch), str(max(1, args.steps // 4))], capture_output=True, text=True, timeout=900)
        try:
            dropin = json.loads(r.stdout.strip().splitlines()[-1])
            dropin["through"] = "the reference's unchanged binding.cpp + headers, Policy::run -> ssb200_homotopy_solve_f32 (dropin/ss_b200_policies.cpp)"
        except Exception:
            dropin = {"error": (r.stderr or r.stdout)[-300:]}
    line = {"metric": "homotopy signals/sec (single-signal calls)", "value": value, "unit": "signals/s", "n_gpus": 1, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * secs / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"C1: homotopy A {M}x{N} float32, k={cfg['k']}, tol {cfg['tol']}, max_iter {cfg['max_iter']}, {batch} signals one call each",
                       "mean_iters": float(its.mean()), "us_per_iteration": 1e6 * secs / (args.steps * batch * (its.mean() + 1))},
            "e2e": {"value": value, "unit": "signals/s", "h2d_bytes_per_step": batch * M * 4, "d2h_bytes_per_step": batch * (N * 4 + 12)},
            "gpu_launches": solver.counters()["launches"],
            "roofline": {"kernel": "homotopy_single_kernel", "bound": "hbm", "achieved": nbytes / secs / 1e9, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                         "frac": nbytes / secs / 1e9 / peaks["hbm_gbs"], "traffic": None,
                         "note": "A (1 MB) is L2-resident: what bounds this path is the latency chain of an iteration (two cluster barriers, one L2 and two DSMEM round trips, the K x K algebra in one CTA) + call overhead, not bandwidth"},
            "dropin": dropin, "cpu_baseline": cpu}
    emit(line)


def run_b200(args, cfg):
    import torch
    import torch.distributed as dist
    import sparsesolvers as ss
    from sparsesolvers import distributed as ssd

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")   # stdout carries the one JSON line only (NCCL prints its version there)
        dist.init_process_group("nccl", device_id=dev)
    assert world == args.gpus, f"--gpus {args.gpus} but WORLD_SIZE={world}"

    total = args.batch or cfg["batch"]
    strong = world > 1 and args.scaling == "strong"
    # strong scaling: the config's batch sharded over the ranks (rank g solves signals [lo, hi), its own seeded draw of them);
    # weak scaling: the config's batch on every rank
    lo, hi = ssd.shard_bounds(total, world, rank) if strong else (0, total)
    batch = hi - lo
    dt = np.dtype(cfg["dtype"])
    tdt = torch.float32 if dt == np.float32 else torch.float64
    M, N = cfg["m"], cfg["n"]

    # the dictionary comes from rank 0 over NCCL; signals are generated where they are solved
    A, Y = make_workload(cfg, batch, seed=4321 + 1000 * rank)
    if world > 1:
        dA = ssd.broadcast_dictionary(torch.from_numpy(A) if rank == 0 else torch.empty((M, N), dtype=tdt), dist, device=dev)
    else:
        dA = torch.from_numpy(A).to(dev)
    torch.cuda.synchronize()

    import ctypes as C
    from sparsesolvers._lib import lib, check
    solver = ss.Homotopy.__new__(ss.Homotopy)            # dictionary already on the device: create_dev
    solver._sfx, solver._ct = ("f32", C.c_float) if dt == np.float32 else ("f64", C.c_double)
    solver.dtype, solver.shape, solver.device, solver._h = dt, (M, N), local, C.c_void_p()
    check(getattr(lib(), f"ssb200_homotopy_create_dev_{solver._sfx}")(
        C.c_void_p(dA.data_ptr()), C.c_size_t(M), C.c_size_t(N), C.c_ssize_t(dA.stride(0)), C.c_ssize_t(dA.stride(1)),
        C.c_int(local), C.byref(solver._h)))
    solver.set_contraction_path(args.path)
    solver.set_residual_mode(args.residual)
    residual_mode, coherence = solver.residual_info()
    stream = torch.cuda.current_stream().cuda_stream
    cap = max(1, min(int(cfg["max_iter"]) + 1, N))       # non-zeros a solution can have: the sparse result's row width

    class Buffers:
        """device buffers of one batch size + the gather targets on rank 0"""
        def __init__(self, Yh):
            n = Yh.shape[0]
            self.n = n
            self.dY = torch.from_numpy(Yh).to(dev)
            self.dX = torch.empty((n, N), dtype=tdt, device=dev)
            self.dIt = torch.empty(n, dtype=torch.int32, device=dev)
            self.dErr = torch.empty(n, dtype=torch.float64, device=dev)
            if world > 1:
                # everything rank 0 is to receive sits in ONE buffer -- sparse rows (index, value) x cap, counts, iterations,
                # solution errors -- and the solver and the compaction kernel write straight into its views: one NCCL gather per step
                up = lambda v: (v + 15) // 16 * 16
                vsz = self.dX.element_size()
                o_val = up(n * cap * 4); o_cnt = o_val + up(n * cap * vsz); o_it = o_cnt + up(n * 4); o_err = o_it + up(n * 4)
                self.pack = torch.zeros(o_err + n * 8, dtype=torch.uint8, device=dev)
                self.idx = self.pack[:n * cap * 4].view(torch.int32).view(n, cap)
                self.val = self.pack[o_val:o_val + n * cap * vsz].view(tdt).view(n, cap)
                self.cnt = self.pack[o_cnt:o_cnt + n * 4].view(torch.int32)
                self.dIt = self.pack[o_it:o_it + n * 4].view(torch.int32)
                self.dErr = self.pack[o_err:].view(torch.float64)
                self.tail = self.pack[o_cnt:]                      # what travels beside a DENSE gather (--gather-x)
                self.o_cnt, self.o_it, self.o_err = o_cnt, o_it, o_err
                mk = (lambda t: [torch.empty_like(t) for _ in range(world)]) if rank == 0 else (lambda t: None)
                self.g_pack = mk(self.tail if args.gather_x else self.pack)
                self.g_X = mk(self.dX) if args.gather_x else None

    def step_dev(B):
        solver.solve_batch_ptr(B.dY.data_ptr(), B.n, M, B.dX.data_ptr(), N, B.dIt.data_ptr(), B.dErr.data_ptr(),
                               cfg["tol"], cfg["max_iter"], True, stream)
        if world > 1:
            # results to rank 0 over NCCL: the coefficients in their sparse form ((index, value) x cap + count per signal: a
            # solution has at most cap = max_iter + 1 non-zeros; 0.65 KB instead of 8 KB per signal at C2) and the reports
            if args.gather_x:
                dist.gather(B.dX, B.g_X, dst=0)
                dist.gather(B.tail, B.g_pack, dst=0)
            else:
                check(getattr(lib(), f"ssb200_compress_rows_dev_{solver._sfx}")(
                    C.c_void_p(B.dX.data_ptr()), C.c_size_t(B.n), C.c_size_t(N), C.c_size_t(N), C.c_size_t(cap),
                    C.c_void_p(B.idx.data_ptr()), C.c_void_p(B.val.data_ptr()), C.c_void_p(B.cnt.data_ptr()), C.c_int(local),
                    C.c_void_p(stream)))
                dist.gather(B.pack, B.g_pack, dst=0)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(B, steps):
        """`steps` steps between one pair of CUDA events, barrier + synchronize on both sides; (ms max over ranks, per-step ms)"""
        barrier()
        marks = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
        marks[0].record()
        for i in range(steps):
            step_dev(B)
            marks[i + 1].record()
        barrier()
        ms = marks[0].elapsed_time(marks[-1])
        per = [marks[i].elapsed_time(marks[i + 1]) for i in range(steps)]
        if world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t[0].item())
        return ms, per

    B = Buffers(Y)
    dY, dX, dIt, dErr = B.dY, B.dX, B.dIt, B.dErr
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    for _ in range(args.warmup):
        step_dev(B)
    # ---- the timed region: K steps, no per-kernel events inside (they cost ~2 ms per step: every record drains the
    # launch pipeline); one event per step boundary only, for the per-step spread
    solver.reset_counters()
    solver.set_timing(False)
    t_wall0 = time.perf_counter()
    ms, step_ms = timed(B, args.steps)
    if rank == 0:
        sampler.window(t_wall0, time.perf_counter())
    launches_timed = solver.counters()["launches"]
    gathered = None
    if world > 1 and rank == 0 and not args.gather_x:      # what rank 0 holds after a step: every rank's reports and sparse rows
        its = torch.cat([t[B.o_it:B.o_it + B.n * 4].view(torch.int32) for t in B.g_pack])
        cnts = torch.cat([t[B.o_cnt:B.o_cnt + B.n * 4].view(torch.int32) for t in B.g_pack])
        gathered = {"signals": int(its.shape[0]), "mean_iters": float(its.double().mean().item()), "nonzeros": int(cnts.sum().item()),
                    "bytes_per_step": int(sum(t.numel() for t in B.g_pack)), "collectives_per_step": 1}
    # ---- the same K steps again with every launch of the lockstep loop bracketed by CUDA events on its stream: the
    # per-kernel durations behind `roofline` / `kernels` (shares are taken against this instrumented pass's own time)
    solver.reset_counters()
    solver.set_timing(not args.no_kernel_timing)
    barrier()
    p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    p0.record()
    for _ in range(args.steps):
        step_dev(B)
    p1.record()
    barrier()
    ms_prof = p0.elapsed_time(p1)
    clocks = sampler.stop() if rank == 0 else None
    solver.set_timing(False)
    counters, timing, prof = solver.counters(), solver.timing(), solver.kernel_profile()
    iters_mean = float(dIt.to(torch.float64).mean().item())
    converged = float((dErr <= cfg["tol"]).to(torch.float64).mean().item())
    it_default = dIt.cpu().numpy().astype(np.uint32)
    X_sample = dX[: min(batch, 1024)].cpu().numpy()
    err_sample = dErr[: min(batch, 1024)].cpu().numpy()

    # ---- the other residual mode on the same inputs (what exactness costs / what the incremental form would give)
    other = None
    if not args.no_exact_extra and dt == np.float32:
        other_mode = "incremental" if residual_mode == "exact" else "exact"
        solver.set_residual_mode(other_mode)
        for _ in range(2):
            step_dev(B)
        k = max(3, min(args.steps, 10))
        ms_o, _ = timed(B, k)
        it_o = dIt.cpu().numpy().astype(np.uint32)
        other = {"mode": other_mode, "ms_per_step": ms_o / k, "value": world * batch * k / (ms_o * 1e-3), "unit": "signals/s",
                 "mean_iters": float(it_o.mean()), "iters_equal_to_default_mode": int((it_o == it_default).sum()), "signals": int(batch)}
        solver.set_residual_mode(args.residual)
        step_dev(B)                                   # leaves the default mode's results in the buffers

    # ---- weak-scaling figure next to the strong one: the config's whole batch on every GPU
    weak = None
    if strong and not args.no_weak:
        Aw, Yw = make_workload(cfg, total, seed=4321 + 1000 * rank)
        Bw = Buffers(Yw)
        for _ in range(2):
            step_dev(Bw)
        k = max(3, min(args.steps, 10))
        ms_w, _ = timed(Bw, k)
        weak = {"value": world * total * k / (ms_w * 1e-3), "unit": "signals/s", "ms_per_step": ms_w / k, "batch_per_gpu": total,
                "scaling": "weak", "steps": k}
        del Bw, Yw

    # ---- end to end: pinned host buffers through the host-pointer entry, copies inside the timed region
    e2e = None
    if not args.no_e2e:
        hY = torch.from_numpy(Y).pin_memory()
        hX = torch.empty((batch, N), dtype=tdt).pin_memory()
        hIt = torch.empty(batch, dtype=torch.int32).pin_memory()
        hErr = torch.empty(batch, dtype=torch.float64).pin_memory()

        def step_host():
            solver.solve_batch_ptr(hY.data_ptr(), batch, M, hX.data_ptr(), N, hIt.data_ptr(), hErr.data_ptr(),
                                   cfg["tol"], cfg["max_iter"], False)
        for _ in range(max(3, args.warmup)):          # warm-up of the host-pointer entry (staging buffers, lanes, helper threads)
            step_host()
        barrier()
        solver.reset_counters()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            step_host()
        barrier()
        e2e_s = time.perf_counter() - t0
        h2d_b, d2h_b = solver.transfer_bytes()        # counted by the library from the copies it issues
        # the same through the SPARSE host entry (extra key): (index, value) x cap + count per signal straight into the caller's
        # arrays -- no dense X, so no zero-fill and no scatter on the host (what bounds the dense entry at 8 GPUs per box)
        hIdx = torch.empty((batch, cap), dtype=torch.int32).pin_memory()
        hVal = torch.empty((batch, cap), dtype=tdt).pin_memory()
        hCnt = torch.empty(batch, dtype=torch.int32).pin_memory()
        sparse_fn = getattr(lib(), f"ssb200_homotopy_solve_batch_sparse_{solver._sfx}")

        def step_host_sparse():
            check(sparse_fn(solver._h, C.c_void_p(hY.data_ptr()), C.c_size_t(batch), C.c_size_t(M), solver._ct(cfg["tol"]), C.c_uint32(cfg["max_iter"]),
                            C.c_void_p(hIdx.data_ptr()), C.c_void_p(hVal.data_ptr()), C.c_void_p(hCnt.data_ptr()), C.c_size_t(cap),
                            C.c_void_p(hIt.data_ptr()), C.c_void_p(hErr.data_ptr())))
        for _ in range(3):
            step_host_sparse()
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            step_host_sparse()
        barrier()
        e2e_sparse_s = time.perf_counter() - t0
        step_host()                                   # the dense result again: hIt below is compared with the device entry's
        # (stragglers are finished by a different kernel depending on the piece they travel in: rounding-level differences there)
        e2e_iter_mismatch = int((hIt != dIt.cpu()).sum().item())
        assert e2e_iter_mismatch <= max(2, batch // 4096), f"host-pointer and device-pointer entries disagree on {e2e_iter_mismatch} iteration counts"
    if world > 1:
        t = torch.tensor([e2e_s if not args.no_e2e else 0.0, e2e_sparse_s if not args.no_e2e else 0.0], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_max, e2e_sparse_max = float(t[0].item()), float(t[1].item())
    else:
        e2e_max, e2e_sparse_max = (e2e_s, e2e_sparse_s) if not args.no_e2e else (0.0, 0.0)
    if not args.no_e2e:
        isz = dt.itemsize
        e2e = {"value": world * batch * args.steps / e2e_max, "unit": "signals/s",
               "h2d_bytes_per_step": world * h2d_b // args.steps, "d2h_bytes_per_step": world * d2h_b // args.steps,
               "result": "dense X (batch x N) in the caller's pinned host buffer; over PCIe the coefficients travel as "
                         "(index, value) x Kmax + count per signal when N >= 8 Kmax, dense otherwise",
               "dense_result_bytes_per_step": world * batch * (N * isz + 4 + 8), "iteration_counts_differing_from_device_entry": e2e_iter_mismatch,
               "sparse_result": {"value": world * batch * args.steps / e2e_sparse_max, "unit": "signals/s",
                                 "note": "the same call with the result taken in its sparse form (ssb200_homotopy_solve_batch_sparse_*): "
                                         "(index, value) x cap + count per signal in the caller's pinned arrays, no dense X on the host"}}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    peaks, peak_src = measured_peaks()
    # roofline of the dominant kernel (contraction): algorithmic flops = 2*M*N per signal-row
    flops = 2.0 * M * N * counters["contraction_rows"]
    achieved = flops / (timing["contraction_ms"] * 1e-3) / 1e12 if timing["contraction_ms"] > 0 else 0.0
    tensor_path = args.path != "simt" and dt == np.float32
    # fp32-equivalent tensor peak of the three-product split: every fp32 product costs three tensor MMAs -- on fp16
    # containers (default) at the bf16/fp16 rate, on tf32 containers (--path tcgen05-tf32) at half of it
    tf32 = args.path == "tcgen05-tf32" or os.environ.get("SSB200_TF32", "0") not in ("", "0")
    peak = peaks["bf16_tflops_sustained"] / (6.0 if tf32 else 3.0)
    roofline = {"kernel": "contract (Q = A^T P)", "bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
                "frac": achieved / peak, "traffic": None,
                "peak_source": f"{peak_src} bf16_tflops_sustained / " + ("6 (TF32 = bf16/2, 3 MMAs per fp32 product)" if tf32 else
                               "3 (3 fp16-rate MMAs per fp32 product)"),
                "launches": counters["contractions"], "avg_launch_ms": timing["contraction_ms"] / max(1, counters["contractions"]),
                "share_of_step": timing["contraction_ms"] / ms_prof, "step_kernel_share": timing["step_ms"] / ms_prof,
                "path": ("tcgen05 3-split, tf32 containers" if tf32 else "tcgen05 3-split, fp16 containers") if tensor_path else "simt fp32"}
    # every kernel class of the lockstep loop against its own bound.  Algorithmic HBM bytes are counted by the library
    # (ssb200_get_kernel_bytes): c read + written, q read when it was streamed, the K x K inverse read + written, the next
    # right-hand side written.  `traffic` = DRAM bytes of one full-chunk launch from the committed ncu --set full capture.
    hbm_peak = peaks["hbm_gbs"]
    traffic = ncu_traffic(args.config)

    def hbm_entry(name, classes, note, bound="hbm"):
        ms_k = sum(prof[c]["ms"] for c in classes)
        if ms_k <= 0:
            return None
        rows_k = sum(prof[c]["rows"] for c in classes)
        bytes_k = sum(prof[c]["algorithmic_bytes"] for c in classes)
        gbs = bytes_k / (ms_k * 1e-3) / 1e9
        tr = [traffic[KERNEL_NAMES[c]] * prof[c]["rows"] for c in classes if KERNEL_NAMES[c] in traffic and prof[c]["rows"]]
        return {"kernel": name, "bound": bound, "achieved": gbs, "peak": hbm_peak, "unit": "GB/s", "frac": gbs / hbm_peak,
                "traffic": sum(tr) / rows_k if len(tr) == len([c for c in classes if prof[c]["rows"]]) and rows_k else None,
                "traffic_note": "DRAM bytes per full-chunk launch (ncu), row-weighted over the variants",
                "peak_source": f"{peak_src} hbm_gbs", "signal_rows": rows_k, "algorithmic_bytes_per_row": bytes_k / max(1, rows_k),
                "share_of_step": ms_k / ms_prof, "algorithmic_bytes": note}
    KERNEL_NAMES = {"scan_stream": "homotopy_scan_kernel (Q streamed)", "scan_gram": "homotopy_scan_kernel (q from Gram rows)",
                    "algebra": "homotopy_algebra_kernel", "step_mono": "homotopy_step_kernel"}
    k_stream = hbm_entry(KERNEL_NAMES["scan_stream"], ["scan_stream"], "3 x Npad x sizeof(T) per signal-iteration: c read, q read, c written")
    k_gram = hbm_entry(KERNEL_NAMES["scan_gram"], ["scan_gram"], "2 x Npad x sizeof(T) per signal-iteration: c read, c written; the K rows of G "
                       "per signal come from L2, which is what bounds this variant (DESIGN.md 3)")
    if k_gram:   # what actually bounds this variant: K rows of G per signal through L2, on top of the HBM stream
        l2_gbs = (prof["scan_gram"]["l2_bytes"] + prof["scan_gram"]["algorithmic_bytes"]) / (prof["scan_gram"]["ms"] * 1e-3) / 1e9
        k_gram["l2"] = {"achieved": l2_gbs, "peak": 12000.0, "unit": "GB/s", "frac": l2_gbs / 12000.0,
                        "peak_source": "this repo's own measurement of L2 -> SM streaming on B200 (TMA operand stream of the contraction with "
                                       "everything else switched off, tools/tc_dbg.py, DESIGN.md 3); not in MEASURED_PEAKS.json",
                        "bytes": "K <= iteration-number rows of G (Npad x sizeof(T) each) per signal + the c stream"}
    k_scan = hbm_entry("homotopy_scan_kernel (both variants)", ["scan_stream", "scan_gram"], "sum of the two variants")
    k_alg = hbm_entry(KERNEL_NAMES["algebra"], ["algebra"], "K x K inverse read + written (K <= iteration number), next right-hand side written; "
                      "the K-atom gather of p = A_G d comes from L2, which is what bounds it", bound="hbm (L2-bound in practice)")
    k_mono = hbm_entry("homotopy_step_kernel (scan + algebra in one CTA per signal: large active sets, small batches)", ["step_mono"],
                       "3 x Npad x sizeof(T) + K x K inverse read + written + right-hand side written, per signal-iteration")
    roofline["traffic"] = traffic.get("contract_x3_pair_kernel")
    kernels = [k for k in (roofline, k_scan, k_stream, k_gram, k_alg, k_mono) if k]
    # `roofline` names the kernel with the largest measured share of the step
    for k in (k_scan, k_mono):
        if k and k["share_of_step"] > roofline["share_of_step"]:
            roofline = dict(k)
    whole = {"achieved": 2.0 * M * N * world * batch * args.steps * (iters_mean + 1) / (ms * 1e-3) / 1e12,
             "unit": "TFLOP/s", "note": "whole-solve algorithmic flops 2*M*N*(iters+1) per signal / step time"}
    whole["frac"] = whole["achieved"] / (peak * world)

    cpu, parity_block = None, None
    if not args.no_cpu_baseline:
        threads = host_threads()
        sample = args.cpu_sample or (512 if args.config == "C2" else 64)
        sample = min(sample, batch)
        secs, kind, it_ref, X_ref, err_ref = cpu_reference_solve(A, Y[:sample], cfg, threads, full=True)
        cpu = {"value": sample / secs, "unit": "signals/s", "cores": threads, "kind": kind,
               "sample": f"first {sample} signals of rank 0's batch, {threads} host threads x 1 BLAS thread, {secs:.1f} s",
               "iters_match_gpu": int((it_ref == it_default[:sample]).sum())}
        # the parity gate's counts on that sample (tests/parity.py): GPU against the CPU path it is timed beside
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import parity as _parity
        c0 = np.abs(Y[:sample].astype(np.float64) @ A.astype(np.float64)).max(axis=1)
        n_par = min(sample, len(X_sample))
        pc = _parity.counts(_parity.compare(X_sample[:n_par], it_default[:n_par], err_sample[:n_par], X_ref[:n_par], it_ref[:n_par], err_ref[:n_par], c0[:n_par]))
        parity_block = {"n": pc["n"], "against": kind, "iter": pc["iter"], "support": pc["support"], "soft_support": pc["soft"],
                        "coef_gt_1e-4": pc["coef"], "coef_max_rel": pc["coef_max"], "solution_error": pc["err"],
                        "note": "mismatch counts of tests/parity.py on the CPU sample; two CPU builds of the reference show the "
                                "same classes of rounding-level mismatches against each other (profiles/r2_parity.json)"}

    line = {
        "metric": "homotopy signals/sec", "value": world * batch * args.steps / (ms * 1e-3), "unit": "signals/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
        "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "f32" if dt == np.float32 else "f64",
        "data": "synthetic",
        "config": {"workload": f"{args.config}: homotopy A {M}x{N} {cfg['dtype']}, {total} k={cfg['k']} signals "
                               + ("sharded over the GPUs" if strong else "per GPU") + f", tol {cfg['tol']}, max_iter {cfg['max_iter']}",
                   "batch_total": world * batch, "batch_per_gpu": batch, "mean_iters": iters_mean, "converged_frac": converged,
                   "l2": "working set (C,Q,X: 3 x %.0f MB per GPU) exceeds the 126 MB L2; no explicit flush" % (batch * N * dt.itemsize / 1e6),
                   "contraction_path": args.path,
                   "gather": None if world == 1 else ("dense X + reports to rank 0 (NCCL gather), inside the timed region" if args.gather_x else
                                                      "sparse coefficients ((index, value) x %d + count per signal) + reports to rank 0 (NCCL gather), inside the timed region" % cap)},
        "residual": {"mode": residual_mode, "requested": args.residual, "mutual_coherence": coherence, "other_mode": other,
                     "note": "exact = c re-formed from r = y - A_G x_G every iteration (homotopy-cpu.cpp:255); incremental = c -= gamma q; auto "
                             "takes exact above coherence 0.7 (include/ssb200.h)"},
        "gathered": gathered, "weak": weak, "parity": parity_block,
        "e2e": e2e, "gpu_launches": launches_timed, "roofline": roofline, "kernels": kernels, "whole_solve": whole,
        "step_ms": {"median": float(np.median(step_ms)), "min": min(step_ms), "max": max(step_ms),
                    "note": "per-step device time inside the timed region (rank 0); `value` uses the whole region"},
        "kernel_timing": {"ms_per_step_instrumented": ms_prof / args.steps,
                          "note": "per-kernel CUDA events are recorded in a second pass of the same K steps right after the "
                                  "timed region; shares are fractions of that pass"},
        "cpu_baseline": cpu, "clocks": clocks,
    }
    emit(line)
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse()
    cfg = CONFIGS[args.config]
    if args.impl == "reference":
        run_reference(args, cfg)
    elif cfg.get("solver") == "irls":
        run_irls(args, cfg)
    elif cfg.get("solver") == "single":
        run_single(args, cfg)
    elif cfg.get("solver") == "grid":
        run_grid(args, cfg)
    else:
        run_b200(args, cfg)


if __name__ == "__main__":
    main()
