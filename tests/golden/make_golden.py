"""Generates tests/golden/*.npz from the UNMODIFIED reference (oracle/_ref, built
from /root/reference by oracle/ref_build/Makefile).  Run in the build container:

    make -C oracle/ref_build && python tests/golden/make_golden.py

The reference ships no golden vectors, so these fixtures ARE the
reference's outputs on seeded inputs: inputs and outputs are stored together and
the file travels to the GPU box, where /root/reference does not exist.
OpenBLAS used by the reference here: scipy wheel 0.3.31.dev (SkylakeX kernels).
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "sparse-solvers_b200"))

from oracle import ref                       # noqa: E402
from sparsesolvers import synth              # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def homotopy_case(name, m, n, k, batch, dtype, tol, max_iter, seed, normalise=True, order="C"):
    A = synth.dictionary(m, n, dtype=dtype, seed=seed)
    if order == "F":
        A = np.asfortranarray(A)
    Y, X0 = synth.sparse_signals(A, batch, k, seed=seed + 1, normalise_sign=normalise)
    X, it, err = ref.homotopy(A, Y, tol, max_iter)
    np.savez_compressed(os.path.join(OUT, name + ".npz"), A=np.ascontiguousarray(A), Y=Y, X0=X0, X=X,
                        iters=it, errs=err, tol=np.float64(tol), max_iter=np.int64(max_iter),
                        blas=np.bytes_(ref.blas_config()))
    print(name, "iters", it.min(), it.mean(), it.max(), "max err", err.max())


def irls_case(name, m, n, k, batch, dtype, tol, max_iter, seed):
    rng = np.random.default_rng(seed)
    A = (rng.standard_normal((m, n)) / np.sqrt(m)).astype(dtype)
    X0 = np.zeros((batch, n), dtype=dtype)
    for b in range(batch):
        idx = rng.choice(n, size=k, replace=False)
        X0[b, idx] = 1.0 + np.abs(rng.standard_normal(k))
    Y = (X0.astype(np.float64) @ A.astype(np.float64).T).astype(dtype)
    X, it, err, spd = ref.irls(A, Y, tol, max_iter)
    np.savez_compressed(os.path.join(OUT, name + ".npz"), A=A, Y=Y, X0=X0, X=X, iters=it, errs=err, spd=spd,
                        tol=np.float64(tol), max_iter=np.int64(max_iter), blas=np.bytes_(ref.blas_config()))
    print(name, "iters", it, "spd", spd.astype(int), "nan", np.isnan(X).any(axis=1).astype(int))


def worked_example():
    """The 5x5 example of the reference's Python prototype (bindings/python/reference/main.py:19-33),
    solved by the reference's C++ solvers."""
    A = np.array([[0.25, 0.25, 0.29, 0.15, 0.14],
                  [0.20, 0.15, 0.02, 0.16, 0.27],
                  [0.15, 0.16, 0.29, 0.07, 0.09],
                  [0.12, 0.25, 0.07, 0.25, 0.28],
                  [0.20, 0.17, 0.29, 0.25, 0.14]], dtype=np.float32)
    b = np.array([[0.27, 0.12, 0.25, 0.02, 0.27]], dtype=np.float32)
    Xh, ith, eh = ref.homotopy(A, b, 0.05, 5)
    Xi, iti, ei, si = ref.irls(A, b, 0.05, 5)
    np.savez_compressed(os.path.join(OUT, "worked_5x5.npz"), A=A, Y=b, Xh=Xh, ith=ith, eh=eh,
                        Xi=Xi, iti=iti, ei=ei, si=si)
    print("worked 5x5 homotopy", Xh, ith, eh, "\n          irls", Xi, iti, ei, si)


if __name__ == "__main__":
    ref.set_blas_threads(1)
    homotopy_case("homotopy_f32_64x256_k4", 64, 256, 4, 48, np.float32, 1e-3, 32, 11)
    homotopy_case("homotopy_f64_64x256_k4", 64, 256, 4, 48, np.float64, 1e-3, 32, 11)
    homotopy_case("homotopy_f32_96x160_k6_colmajor", 96, 160, 6, 24, np.float32, 1e-3, 60, 23, order="F")
    homotopy_case("homotopy_f32_256x1024_k16", 256, 1024, 16, 16, np.float32, 1e-3, 48, 1234)
    # un-normalised signs: the sign(|c|) init quirk (homotopy-cpu.cpp:223-227); short horizon so the
    # trajectory is still rounding-stable (DESIGN.md "noise floor")
    homotopy_case("homotopy_f64_quirk_64x256_k8", 64, 256, 8, 32, np.float64, 1e-3, 6, 31, normalise=False)
    irls_case("irls_f64_60x16_k2", 60, 16, 2, 8, np.float64, 1e-3, 50, 5)
    irls_case("irls_f64_128x32_k1", 128, 32, 1, 8, np.float64, 1e-3, 50, 7)
    irls_case("irls_f32_60x16_k2", 60, 16, 2, 8, np.float32, 1e-3, 50, 6)
    worked_example()
