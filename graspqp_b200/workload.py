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


# BASELINE.json's multi-GPU configs as GLOBAL object lists, dealt round-robin over the ranks (shard.objects_of_rank):
# config 3 = 64 meshes (8 per GPU on 8 GPUs), config 4 = 256 meshes (32 per GPU), config 5 = 32 meshes at 2/4/8 GPUs
GLOBAL_OBJECTS = {"config3": 64, "config4": 256, "config4_robotiq2": 256, "config5": 32}


def build(name="config2", rank=0, n_surface=2500, batch_each=None, n_objects=None, seed=1, world=None):
    """-> dict(pack, face_verts [list of (F,3,3)], surface (O,P_s,3), pose (O*B,D), idx (O*B,n_c) int32, batch_each)

    world=None: every rank builds its own copy of the config's per-GPU object list (weak scaling; rank only changes the
    seeds).  world=W: the config's GLOBAL object list (GLOBAL_OBJECTS, or n_objects) is dealt round-robin over W ranks
    and this rank builds its share (the reference's job-level partition, scripts/fit_all.sh:154-248)."""
    hand, kinds, nf, B, n_c = CONFIGS[name]
    obj_seed = lambda i: seed + 100 * rank + i
    if world is not None:
        from . import shard
        n_global = n_objects if n_objects is not None else GLOBAL_OBJECTS.get(name, len(kinds) * world)
        mine = shard.objects_of_rank(n_global, rank, world)
        all_kinds = kinds
        kinds = [all_kinds[g % len(all_kinds)] for g in mine]
        obj_seed = lambda i: seed + mine[i]
    elif n_objects is not None:
        kinds = (kinds * ((n_objects + len(kinds) - 1) // len(kinds)))[:n_objects]
    if batch_each is not None:
        B = batch_each
    pack = hand_pack.load_pack(hand)
    fvs, verts, surf = [], [], []
    for i, k in enumerate(kinds):
        s = obj_seed(i)
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
