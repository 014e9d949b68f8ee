"""Synthetic workloads of BASELINE.json's configs (the reference ships no dataset, scripts/fit.py:124-125)."""
import numpy as np

from . import hand_pack, synth

CONFIGS = {
    # name: (hand, mesh kinds, faces per mesh, batch per object, n_contact)
    "config1": ("allegro", ["sphere"], 2000, 32, 12),
    "config2": ("allegro", ["box", "cylinder", "torus", "hull", "hull", "hull", "hull", "hull"], 10000, 1024, 12),
    "headline": ("allegro", ["box", "cylinder", "torus", "hull", "hull", "hull", "hull", "hull"], 10000, 4096, 12),
    # per-GPU shares of the 8-GPU configs (objects are dealt round-robin over the ranks, shard.py)
    "config3": ("shadow_hand", ["box", "cylinder", "torus", "hull", "hull", "hull", "hull", "hull"], 10000, 4096, 12),
    "config4": ("panda", ["box", "cylinder", "torus", "hull"] * 8, 10000, 16384, 4),
    "config4_robotiq2": ("robotiq2", ["box", "cylinder", "torus", "hull"] * 8, 10000, 16384, 4),
    "config5": ("allegro", ["hull"] * 4, 50000, 4096, 16),
}


def build(name="config2", rank=0, n_surface=2500, batch_each=None, n_objects=None, seed=1):
    """-> dict(pack, face_verts [list of (F,3,3)], surface (O,P_s,3), pose (O*B,D), idx (O*B,n_c) int32, batch_each)"""
    hand, kinds, nf, B, n_c = CONFIGS[name]
    if n_objects is not None:
        kinds = (kinds * ((n_objects + len(kinds) - 1) // len(kinds)))[:n_objects]
    if batch_each is not None:
        B = batch_each
    pack = hand_pack.load_pack(hand)
    fvs, verts, surf = [], [], []
    for i, k in enumerate(kinds):
        s = seed + 100 * rank + i
        v, f = synth.make_mesh(k, nf, s)
        verts.append(v)
        fvs.append(synth.face_vertices(v, f))
        surf.append(synth.sample_surface(v, f, n_surface, seed=s, oversample=2))
    pose, idx = synth.random_grasp_init(pack, verts, B, n_c, seed + 100 * rank)
    return dict(name=name, hand=hand, pack=pack, face_verts=fvs, surface=np.stack(surf), pose=pose,
                idx=idx.astype(np.int32), batch_each=B, n_contact=n_c, kinds=kinds, faces_per_mesh=nf)


def bfe_flops_per_grasp_step(pack, n_c, faces_obj, n_surface, c_pt=80.0):
    """brute-force-equivalent FP32 work of one grasp-step."""
    f_hand = int(pack["face_offset"][-1])
    return c_pt * (n_c * faces_obj + n_surface * f_hand) + 0.65e6 + 0.05e6
