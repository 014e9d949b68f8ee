"""Diagnostic: at the trajectory's outlier steps, separate input differences (contacts, normals) from QP differences."""
import os, sys
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
import numpy as np, torch
from graspqp_b200 import _lib, engine, hand_pack, synth
from oracle.pipeline import FitLoopOracle
from oracle.hand import HandOracle
from oracle import qp as oqp
lib, be = _lib.lib(), engine.TorchBackend("cuda:0")
T, B, n_pts, n_c, seed = 500, 6, 128, 12, 21
pack = hand_pack.load_pack("allegro")
v, f = synth.make_mesh("hull", 400, seed); fv = synth.face_vertices(v, f); sp = synth.sample_surface(v, f, n_pts, seed=seed)
pose, idx = synth.random_grasp_init(pack, [v], B, n_c, seed); pose[:, :3] *= 0.75
eng = engine.FitEngine(lib, be, pack, [fv], sp[None], B, pose, idx.astype(np.int32))
orc = FitLoopOracle(pack, [fv], [sp], B, pose, idx, dtype=torch.float64)
H = be.to_host
rng = np.random.default_rng(seed)
for t in range(T):
    orc.load_state(H(eng.pose), H(eng.idx), H(eng.grad), H(eng.energy), H(eng.ema), H(eng.step_count))
    us, ir, ua = rng.random((B, n_c), dtype=np.float32), rng.integers(0, len(pack["cand"]), (B, n_c)).astype(np.int32), rng.random(B, dtype=np.float32)
    eng.step(be.from_host(us), be.from_host(ir), be.from_host(ua))
    out = orc.iterate(torch.tensor(us, dtype=torch.float64), torch.tensor(ir).long(), torch.tensor(ua, dtype=torch.float64))
    rel = np.abs(H(eng.new_energy) - out["new_energy"].numpy()) / np.abs(out["new_energy"].numpy())
    for b in np.nonzero(rel > 2e-4)[0]:
        h = HandOracle(pack, torch.float64); h.set_parameters(out["proposal"], out["idx_prop"])
        d, n = orc.obj.cal_distance(h.contact_points)
        print(f"--- step {t} cand {b} rel {rel[b]:.2e} engine stop {int(eng.qp_stop.item())}")
        print(" cpts max diff", np.abs(H(eng.cpts)[b] - h.contact_points[b].numpy()).max())
        print(" obj_dist eng", H(eng.obj_dist)[b]); print(" obj_dist orc", d[b].numpy())
        print(" nrm diff per contact", np.abs(H(eng.obj_nrm)[b] - n[b].numpy()).max(-1))
        # QP on the ENGINE's inputs in the fp64 oracle, whole batch (batch-wide exit)
        tr = []
        cp, nn = torch.tensor(H(eng.cpts), dtype=torch.float64), torch.tensor(H(eng.obj_nrm), dtype=torch.float64)
        E, lam, info = oqp.graspqp_energy(cp, nn, orc.obj.cog)
        print(" E_fc engine", H(eng.losses_new)[1], "\n E_fc oracle(engine inputs)", E.numpy(), "\n E_fc oracle", out["new_losses"]["E_fc"].numpy())
        F = info["F"]; nq = F.shape[-1]
        Q = F.mT @ F + torch.eye(nq, dtype=torch.float64) * 1e-4
        G = torch.cat([torch.eye(nq), -torch.eye(nq)], 0).double().unsqueeze(0).expand(B, -1, -1).contiguous()
        hh = torch.cat([torch.full((B, nq), 21.0), torch.full((B, nq), -1.0)], -1).double()
        oqp.pdipm_forward(Q, torch.zeros(B, nq, dtype=torch.float64), G, hh, trace=tr)
        print(" oracle iterations", len(tr), "last resid", tr[-1].numpy())
