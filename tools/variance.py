"""Per-step device time of the C2 batched solve (device-resident buffers), step by step, over handle re-creations:
   is the run-to-run spread per process, per handle or per step?   python tools/variance.py [handles] [steps]"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "sparse-solvers_b200"))
import numpy as np, torch
import sparsesolvers as ss
from sparsesolvers import synth
handles = int(sys.argv[1]) if len(sys.argv) > 1 else 3
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 8
A = synth.dictionary(512, 2048, seed=1236)
Y, _ = synth.sparse_signals(A, 65536, 32, seed=4321)
dev = torch.device("cuda", 0)
dY = torch.from_numpy(Y).to(dev)
dX = torch.empty((65536, 2048), dtype=torch.float32, device=dev)
dIt = torch.empty(65536, dtype=torch.int32, device=dev); dErr = torch.empty(65536, dtype=torch.float64, device=dev)
st = torch.cuda.current_stream().cuda_stream
for h in range(handles):
    s = ss.Homotopy(A)
    ms, wall = [], []
    for i in range(steps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); t0 = time.perf_counter(); e0.record()
        s.solve_batch_ptr(dY.data_ptr(), 65536, 512, dX.data_ptr(), 2048, dIt.data_ptr(), dErr.data_ptr(), 1e-3, 80, True, st)
        e1.record(); torch.cuda.synchronize(); wall.append(1e3 * (time.perf_counter() - t0)); ms.append(e0.elapsed_time(e1))
    print(f"handle {h}: device ms", " ".join(f"{m:.1f}" for m in ms), "| wall", " ".join(f"{m:.1f}" for m in wall), flush=True)
    del s
