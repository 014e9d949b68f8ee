"""Per-stage device times of one grasp-step (CUDA events, each stage launched alone `reps` times on the state of a
warmed-up engine).  python tools/bench_kernels.py [workload] [reps]"""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from graspqp_b200 import _lib, engine, workload

name = sys.argv[1] if len(sys.argv) > 1 else "config2"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 20
lib, be = _lib.lib(), engine.TorchBackend("cuda:0")
w = workload.build(name)
e = engine.FitEngine(lib, be, w["pack"], w["face_verts"], w["surface"], w["batch_each"], w["pose"], w["idx"])
for _ in range(30):
    e.step()
Bt, n_c, O = e.Bt, e.n_c, e.n_obj
stages = dict(e.stage_calls())
stages["full step"] = lambda: e.step() or 0
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
print(f"{name}: Bt={Bt} n_c={n_c} objects={O}")
for k, fn in stages.items():
    for _ in range(3):
        assert fn() == 0, lib.gqp_last_error()
    ts = []
    for _ in range(reps):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    print(f"  {k:22s} {np.median(ts)*1e3:9.1f} us  (min {min(ts)*1e3:.1f})")
