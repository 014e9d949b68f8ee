import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from graspqp_b200 import _lib, engine, workload
lib, be = _lib.lib(), engine.TorchBackend("cuda:0")
for name in ["config2", "config3"]:
    w = workload.build(name)
    e = engine.FitEngine(lib, be, w["pack"], w["face_verts"], w["surface"], w["batch_each"], w["pose"], w["idx"])
    for _ in range(30): e.step()
    e.stage_calls()["hand_distance"]()
    torch.cuda.synchronize()
    Bt, n_pts, O = e.Bt, e.n_pts, e.n_obj
    n_chunks = (n_pts + 31) // 32
    off = Bt * n_chunks + O * (n_chunks + 1) * 4 + Bt
    sc = e.pen_scratch.view(torch.int32)
    base = e.pen_scratch.data_ptr() + 4 * off
    if base % 8: off += 1
    cnt = sc[off:off + 4].cpu().numpy()
    q = sc[off + 4 + 2 * Bt: off + 4 + 2 * Bt + 2 * int(cnt[0])].cpu().numpy().reshape(-1, 2)
    nl = np.array([bin(int(v) & 0xffffffff).count("1") for v in q[:20000, 1]])
    dist = torch.empty((Bt, n_pts), device="cuda")
    lib.gqp_hand_distance(e.hand, be.ptr(e.link_tf), be.ptr(e.surf), O, e.batch_each, n_pts, 0, be.ptr(dist), None, be.ptr(e.pen_scratch), 0.0, None, be.stream)
    torch.cuda.synchronize()
    print(name, "Bt", Bt, "groups", cnt[0], "items", cnt[1], "of", Bt * n_chunks, "links/group", nl.mean(), "penetrating samples", int((dist > 0).sum()),
          "within 1mm", int((dist > -0.001).sum()), "within 2mm", int((dist > -0.002).sum()))
