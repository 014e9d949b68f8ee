"""Diagnostic: distribution of qpth's batch-wide exit iteration (fc_stop_iter) over a run of a workload."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from graspqp_b200 import _lib, engine, workload
name = sys.argv[1] if len(sys.argv) > 1 else "config2"
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 60
lib, be = _lib.lib(), engine.TorchBackend("cuda:0")
w = workload.build(name)
eng = engine.FitEngine(lib, be, w["pack"], w["face_verts"], w["surface"], w["batch_each"], w["pose"], w["idx"])
stops, acc = [], []
for t in range(steps):
    eng.step()
    stops.append(int(eng.qp_stop.item())); acc.append(float(eng.accept.float().mean().item()))
print("stops", stops)
print("accept rate", np.round(acc, 2).tolist())
tr = None
print("mean energy", float(eng.energy.mean().item()))
