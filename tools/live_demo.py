"""Closed-loop acquisition demo: the reference's object graph (read threads, ccontrol, sender) free-running
against the simulated dongles of oracle/shim/shim_impl.cc; DSP thread = the drop-in GPU path
(default) or the reference's own CPU code (--cpu).  Prints lag / synced / true delay every 50 ms."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import oracle as O
import numpy as np
t0=time.time()
lv = O.RefLive([120, -35, 4], fs=1024000, phases=[0.3,-1.0,2.0], gains=[0.9,0.8,0.7], lib_path=None if "--cpu" in sys.argv else O.DROPIN_LIB_PATH)
print("created", time.time()-t0, flush=True)
for k in range(60):
    time.sleep(0.05); p = lv.poll()
    print(f"{time.time()-t0:5.2f} lag={p['lag'][1:]} sync={p['synced'][1:]} d={np.round(p['delay'][1:],1)} corr={p['corr'][1:]} blk={p['blocks']} pk={p['packets']}", flush=True)
    if p['synced'].all(): break
lv.close(); print("closed", time.time()-t0)
