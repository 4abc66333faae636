"""Host-pointer entry timing probe (pinned buffers), C2 shape."""
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
for label, env in (("default", {}), ("dense copy-out", {"SSB200_HOST_SPARSE": "0"}), ("nocopy", {"SSB200_HOST_NOCOPY": "1"}), ("1 piece", {"SSB200_HOST_PIECES": "1"}),
                   ("2 pieces", {"SSB200_HOST_PIECES": "2"}), ("3 pieces", {"SSB200_HOST_PIECES": "3"}), ("4 pieces", {"SSB200_HOST_PIECES": "4"}),
                   ("8 pieces", {"SSB200_HOST_PIECES": "8"}), ("4 zero threads", {"SSB200_ZERO_THREADS": "4"}), ("16 zero threads", {"SSB200_ZERO_THREADS": "16"}),
                   ("default again", {})):
    for k in ("SSB200_HOST_NOCOPY", "SSB200_HOST_PIECES", "SSB200_HOST_SPARSE", "SSB200_ZERO_THREADS"):
        os.environ.pop(k, None)
    os.environ.update(env)
    run(); run()
    ts = []
    for _ in range(8):
        t0 = time.perf_counter(); run(); ts.append((time.perf_counter() - t0) * 1e3)
    print(label, "ms/step median %.2f min %.2f max %.2f" % (np.median(ts), min(ts), max(ts)), flush=True)
d = torch.empty((65536, 2048), device="cuda")
torch.cuda.synchronize(); t0 = time.perf_counter(); hX.copy_(d, non_blocking=True); torch.cuda.synchronize()
print("D2H 537 MB ms %.2f" % ((time.perf_counter() - t0) * 1e3))
dy = torch.empty((65536, 512), device="cuda")
torch.cuda.synchronize(); t0 = time.perf_counter(); dy.copy_(hY, non_blocking=True); torch.cuda.synchronize()
print("H2D 134 MB ms %.2f" % ((time.perf_counter() - t0) * 1e3))
