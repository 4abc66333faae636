"""Times the contraction kernels alone (Q = A^T P) and reports accuracy against float64.
    python tools/contract_bench.py [M N rows]"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "sparse-solvers_b200"))
import sparsesolvers as ss  # noqa: E402

M, N, rows = (int(v) for v in sys.argv[1:4]) if len(sys.argv) >= 4 else (512, 2048, 65536)
rng = np.random.default_rng(0)
A = (rng.standard_normal((M, N)) / np.sqrt(M)).astype(np.float32)
P = rng.standard_normal((rows, M)).astype(np.float32)
s = ss.Homotopy(A)
out = {"M": M, "N": N, "rows": rows}
sub = slice(0, min(rows, 2048))
ref = P[sub].astype(np.float64) @ A.astype(np.float64)
scale = np.abs(P[sub].astype(np.float64)) @ np.abs(A.astype(np.float64))
for path in ("tcgen05", "simt"):
    Q, ms = s.debug_contract(P, path, reps=10)
    err = np.abs(Q[sub] - ref)
    out[path] = {"ms": ms, "tflops": 2.0 * M * N * rows / (ms * 1e-3) / 1e12,
                 "max_err_over_sum_abs": float((err / scale).max()),
                 "rms_err_over_rms_q": float(np.sqrt((err ** 2).mean()) / np.sqrt((ref ** 2).mean())),
                 "mean_signed_err_over_rms_q": float(((Q[sub] - ref) * np.sign(ref)).mean() / np.sqrt((ref ** 2).mean()))}
# accumulation-only error: inputs exactly representable in TF32, so every product is exact
At = A.view(np.uint32) & np.uint32(0xFFFFE000)
Pt = P.view(np.uint32) & np.uint32(0xFFFFE000)
A2, P2 = At.view(np.float32), Pt.view(np.float32)
s2 = ss.Homotopy(A2)
ref2 = P2[sub].astype(np.float64) @ A2.astype(np.float64)
for path in ("tcgen05", "simt"):
    Q, _ = s2.debug_contract(P2[sub], path)
    e = Q - ref2
    out[path]["tf32_exact_inputs_rms_err_over_rms_q"] = float(np.sqrt((e ** 2).mean()) / np.sqrt((ref2 ** 2).mean()))
    out[path]["tf32_exact_inputs_signed_bias_over_rms_q"] = float((e * np.sign(ref2)).mean() / np.sqrt((ref2 ** 2).mean()))
print(json.dumps(out))
