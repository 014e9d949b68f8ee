"""Step time along a fit: ms per grasp-step in windows of `win` iterations, plus per-stage times at chosen points.
python tools/bench_trajectory.py [workload] [n_iter] [win]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from graspqp_b200 import _lib, engine, workload

name = sys.argv[1] if len(sys.argv) > 1 else "config2"
n_iter = int(sys.argv[2]) if len(sys.argv) > 2 else 1000
win = int(sys.argv[3]) if len(sys.argv) > 3 else 100
lib, be = _lib.lib(), engine.TorchBackend("cuda:0")
w = workload.build(name)
e = engine.FitEngine(lib, be, w["pack"], w["face_verts"], w["surface"], w["batch_each"], w["pose"], w["idx"])
e.capture()
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def stages():
    out = {}
    for k, fn in e.stage_calls().items():
        ts = []
        for _ in range(7):
            flush.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); fn(); b.record()
            torch.cuda.synchronize()
            ts.append(a.elapsed_time(b))
        out[k] = round(float(np.median(ts)), 3)
    return out


print(f"{name}: Bt={e.Bt}")
for s0 in range(0, n_iter, win):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    a.record()
    for _ in range(win):
        e.step()
    b.record()
    torch.cuda.synchronize()
    en = e.energy.float()
    print(f"iter {s0:5d}-{s0 + win:5d}: {a.elapsed_time(b) / win:7.3f} ms/step  mean E {en.mean().item():9.3f}  "
          f"E_pen {e.losses[engine.TERMS.index('E_pen')].mean().item() if hasattr(e, 'losses') else float('nan'):.4f}")
    if s0 in (0, n_iter - win) or (s0 // win) % 4 == 3:
        print("   stages:", stages())
