"""Where an iteration of the single-signal kernel goes: device time per solve (CUDA events around the launch) against wall time per
call, and -- with the variant build `tools/build_variant.sh prof "-DSSB_SINGLE_PROF=1"`, SSB200_LIB=build/variants/prof/libssb200.so --
clock64 cycles per phase as CTA 0 / thread 0 sees them.   python tools/single_prof.py [M N k max_iter]"""
import ctypes as C, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "sparse-solvers_b200"))
import sparsesolvers as ss
from sparsesolvers import synth
from sparsesolvers._lib import lib
M, N, k, mi = (int(v) for v in sys.argv[1:5]) if len(sys.argv) >= 5 else (256, 1024, 16, 48)
A = synth.dictionary(M, N, dtype=np.float32, seed=1235)
Y, _ = synth.sparse_signals(A, 256, k, seed=7)
s = ss.Homotopy(A)
for b in range(8): s.solve(Y[b], 1e-3, mi)
PH = ["p copy", "correlate", "search+argmin+push", "barrier 1", "partials+argmin", "c update+push", "barrier 2", "lambda", "x,slot,c_G gather",
      "inverse update", "direction", "rhs gather", "barrier 3"]
have = hasattr(lib(), "ssb200_debug_single_profile")
out = (C.c_uint64 * 32)()
if have: lib().ssb200_debug_single_profile(out)
s.set_timing(True); s.reset_counters()
t0 = time.perf_counter(); its = []
for b in range(256):
    x, rep = s.solve(Y[b], 1e-3, mi); its.append(rep.iter)
wall = (time.perf_counter() - t0) / 256
dev = s.timing()["step_ms"] / 256
n = np.mean(its)
print(f"{M}x{N} k={k}: wall {wall * 1e6:.1f} us/solve, kernel {dev * 1e3:.1f} us/solve, mean iters {n:.2f}: kernel {dev * 1e3 / (n + 1):.2f} us/iteration, call overhead {wall * 1e6 - dev * 1e3:.1f} us",
      {e: os.environ[e] for e in os.environ if e.startswith("SSB200_")})
if have:
    lib().ssb200_debug_single_profile(out)
    iters = max(1, out[15])
    tot = sum(out[i] for i in range(13))
    print(f"  cycles per loop iteration as thread 0 of CTA 0 sees them (a probe costs ~60): phases {tot / iters:.0f} + inner slots {sum(out[i] for i in range(16, 24)) / iters:.0f}")
    for i, name in enumerate(PH):
        print(f"    {name:22s} {out[i] / iters:8.0f}")
    SUB = {16: "search: candidates", 17: "search: warp argmin", 18: "search: push", 19: "inverse: u1 + barrier", 20: "inverse: matvec", 21: "inverse: uu, 1/x",
           22: "inverse: K x K update", 23: "gram q: loads + sums"}
    for i, name in SUB.items():
        print(f"      [{name:22s} {out[i] / iters:8.0f}]  (its own slot, not part of the phase it sits in)")
