"""Kernel timeline (torch profiler / CUPTI) of one stage call, cold and warm L2.  python tools/stage_timeline.py [stage]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from torch.profiler import profile, ProfilerActivity
from graspqp_b200 import _lib, engine, workload
lib, be = _lib.lib(), engine.TorchBackend("cuda:0")
w = workload.build("config2")
e = engine.FitEngine(lib, be, w["pack"], w["face_verts"], w["surface"], w["batch_each"], w["pose"], w["idx"])
for _ in range(30): e.step()
fn = e.stage_calls()[sys.argv[1] if len(sys.argv) > 1 else "hand_distance"]
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for _ in range(3): fn()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(3):
        flush.zero_()
        fn()
        torch.cuda.synchronize()
    for _ in range(3):
        fn()
        torch.cuda.synchronize()
ev = [x for x in prof.events() if x.device_type == torch.autograd.DeviceType.CUDA]
ev.sort(key=lambda x: x.time_range.start)
t0 = ev[0].time_range.start
for x in ev:
    print(f"{x.time_range.start - t0:10.1f} us  dur {x.time_range.end - x.time_range.start:8.1f}  {x.name[:60]}")
