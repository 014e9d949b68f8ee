#!/usr/bin/env python
"""Diagnostic (GPU): find the samples whose penetration depth the culled path of gqp_hand_distance drops or adds
relative to the un-culled path, and print the oracle's view of them.  python tools/debug_pen_cull.py shadow_hand"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import cases  # noqa: E402
from graspqp_b200 import _lib, engine, hand_pack, synth  # noqa: E402
from oracle.hand import HandOracle  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "shadow_hand"
B, n_pts, seed = 192, 2500, 5
lib, be = _lib.lib(), engine.TorchBackend("cuda:0")
pack = hand_pack.load_pack(name)
hand = hand_pack.create_hand(lib, pack)
v, f = synth.make_mesh("hull", 2000, seed)
sp = synth.sample_surface(v, f, n_pts, seed=seed, oversample=2)
pose, idx = synth.random_grasp_init(pack, [v], B, 12, seed)
pose[:, :3] *= np.linspace(0.0, 1.0, B, dtype=np.float32)[:, None]
dpose, tf = cases.fk_state(lib, be, hand, pack, pose)
L = len(pack["mesh_frame"])


def run(points, full):
    n = points.shape[0]
    x = be.from_host(np.ascontiguousarray(points[None]), np.float32)
    dist, e = be.empty((B, n), np.float32), be.empty((B,), np.float32)
    sc = be.empty((int(lib.gqp_hand_distance_scratch_floats(1, B, n, 0)),), np.float32)
    _lib.check(lib.gqp_hand_distance(hand, be.ptr(tf), be.ptr(x), 1, B, n, 0, be.ptr(dist) if full else None, be.ptr(e),
                                     be.ptr(sc), 100.0, None, be.stream), lib)
    return be.to_host(e), (be.to_host(dist) if full else None)


e1, d1 = run(sp, True)
e2, _ = run(sp, False)
bad = np.where(np.abs(e1 - e2) > 1e-6 * np.abs(e1) + 1e-9)[0]
print("mismatching candidates", bad, e1[bad], e2[bad])
o = HandOracle(pack, torch.float32)
for b in bad[:3]:
    o.set_parameters(torch.tensor(pose[b:b + 1]), torch.tensor(idx[b:b + 1]).long())
    want = o.cal_distance(torch.tensor(sp).unsqueeze(0)).numpy()[0]
    print(f"candidate {b}: oracle E_pen {np.maximum(want, 0).sum():.6f} full {e1[b]:.6f} culled {e2[b]:.6f}; "
          f"max |full dist - oracle| {np.abs(d1[b] - want).max():.2e}")
    pos = np.where(d1[b] > 0)[0]
    for s in pos:
        es, _ = run(sp[s:s + 1], False)
        if abs(es[b] - d1[b, s]) > 1e-6:
            # per-link oracle distances of this sample
            per_link = []
            for li in range(L):
                T = o.link_matrix(li)[0].numpy()
                R, t = T[:3, :3], T[:3, 3]
                Rg, tg = o.global_rotation[0].numpy(), o.global_translation[0].numpy()
                xl = ((sp[s] - tg) @ Rg - t) @ R
                per_link.append(xl)
            print(f"  sample {s}: full {d1[b, s]:.6f} oracle {want[s]:.6f} culled-alone {es[b]:.6f}")
            from oracle import sdf as osdf
            for li in range(L):
                fo = pack["face_offset"]
                tr = pack["face_verts"][fo[li]:fo[li + 1]]
                if len(tr) == 0:
                    continue
                d2, sg, _, _ = osdf.compute_sdf_port(per_link[li][None].astype(np.float32), tr)[:4]
                lo, hi = tr.reshape(-1, 3).min(0), tr.reshape(-1, 3).max(0)
                inside_box = bool(((per_link[li] >= lo) & (per_link[li] <= hi)).all())
                if sg[0] < 0 or inside_box:
                    print(f"    link {li} {pack['mesh_names'][li]}: d {np.sqrt(d2[0]):.6f} sign {sg[0]} in box {inside_box} "
                          f"local {per_link[li]} box margin {np.minimum(per_link[li] - lo, hi - per_link[li]).min():.2e}")
