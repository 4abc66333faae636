import sys,os; sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/sparse-solvers_b200')
import numpy as np, sparsesolvers as ss
from sparsesolvers import synth
A=synth.dictionary(512,2048,seed=1236); Y,_=synth.sparse_signals(A,65536,32,seed=1)
s=ss.Homotopy(A); mi=int(os.environ.get("MI","20"))
s.solve_batch(Y,1e-3,mi); s.reset_counters(); s.set_timing(True)
s.solve_batch(Y,1e-3,mi); t=s.timing(); c=s.counters()
print(os.environ.get("SSB200_DBG"), "iters",c["iterations"],"step ms/launch", t["step_ms"]/(c["iterations"]+1), "contract ms/launch", t["contraction_ms"]/c["contractions"])
