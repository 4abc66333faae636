"""Host-pointer entry (C2, pinned buffers): piece fractions dealt to the two lanes.  python tools/taper_probe.py"""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "sparse-solvers_b200"))
import torch
import sparsesolvers as ss
from sparsesolvers import synth
A = synth.dictionary(512, 2048, dtype=np.float32, seed=1236)
Y, _ = synth.sparse_signals(A, 65536, 32, seed=4321)
s = ss.Homotopy(A)
hY = torch.from_numpy(Y).pin_memory(); hX = torch.empty((65536, 2048)).pin_memory()
hIt = torch.empty(65536, dtype=torch.int32).pin_memory(); hErr = torch.empty(65536, dtype=torch.float64).pin_memory()
def run():
    s.solve_batch_ptr(hY.data_ptr(), 65536, 512, hX.data_ptr(), 2048, hIt.data_ptr(), hErr.data_ptr(), 1e-3, 80, False)
for t in sys.argv[1:] or ["0.25,0.25,0.1875,0.1875,0.0625,0.0625"]:
    os.environ["SSB200_TAPER"] = t
    run(); run()
    ts = []
    for _ in range(8):
        t0 = time.perf_counter(); run(); ts.append((time.perf_counter() - t0) * 1e3)
    print(t, "ms/step median %.2f min %.2f max %.2f" % (np.median(ts), min(ts), max(ts)), flush=True)
