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
p, st = be.ptr, be.stream
Bt, n_c, O, B = e.Bt, e.n_c, e.n_obj, e.batch_each
row = lambda k: p(e.losses_new) + 4 * Bt * engine.TERMS.index(k)
stages = {
    "fk_forward": lambda: lib.gqp_fk_forward(e.hand, p(e.pose_new), Bt, p(e.link_tf), st),
    "contacts_forward": lambda: lib.gqp_contacts_forward(e.hand, p(e.link_tf), p(e.idx_new), Bt, n_c, p(e.cpts), p(e.cnrm), st),
    "object_distance": lambda: lib.gqp_object_distance_forward(e.objset, B, n_c, p(e.cpts), 1.0, p(e.obj_dist), p(e.obj_nrm), p(e.obj_clst), p(e.obj_sign), st),
    "fc_solve": lambda: lib.gqp_fc_solve(ctypes.byref(e.fc), p(e.cpts), p(e.obj_nrm), p(e.cog), O, B, n_c, p(e.qp_ws), st),
    "fc_finish": lambda: lib.gqp_fc_finish(ctypes.byref(e.fc), p(e.cpts), p(e.cnrm), p(e.obj_dist), p(e.obj_nrm), p(e.obj_dgrad), p(e.cog), O, B, n_c, p(e.qp_ws), 1, row("E_fc"), row("E_dis"), p(e.lam), p(e.g_cpts), p(e.g_cnrm), p(e.qp_stop), st),
    "hand_distance(pen)": lambda: lib.gqp_hand_distance(e.hand, p(e.link_tf), p(e.surf), O, B, e.n_pts, 0, None, row("E_pen"), p(e.pen_scratch), 100.0, p(e.g_link_tf), st),
    "self_pen_joints": lambda: lib.gqp_self_pen_joints(e.hand, p(e.link_tf), p(e.pose_new), Bt, row("E_spen"), row("E_joints"), 10.0, 1.0, p(e.g_link_tf), p(e.g_pose_joints), st),
    "fk_backward": lambda: lib.gqp_fk_backward(e.hand, p(e.pose_new), p(e.g_link_tf), p(e.g_pose_joints), Bt, p(e.grad_new), st),
    "full step": lambda: e.step() or 0,
}
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
