"""Hand packs: the reference's hand description (URDF tree, link meshes, contact candidates, penetration spheres),
compiled once by tools/build_hand_pack.py into graspqp_b200/assets/<hand>.npz, and its marshalling into GqpHandDesc."""
import ctypes
import os

import numpy as np

from . import _lib

ASSETS = os.path.join(os.path.dirname(os.path.abspath(__file__)), "assets")


def load_pack(name):
    path = os.path.join(ASSETS, name + ".npz")
    if not os.path.exists(path):
        raise FileNotFoundError(f"no hand pack {path}; available: {sorted(os.listdir(ASSETS))}")
    z = np.load(path)
    return {k: z[k] for k in z.files}


def make_desc(pack):
    """-> (GqpHandDesc, keepalive list).  All arrays are host numpy buffers."""
    keep = []

    def arr(key, dtype):
        a = np.ascontiguousarray(pack[key], dtype)
        keep.append(a)
        return a.ctypes.data

    d = _lib.GqpHandDesc()
    d.n_frames = len(pack["frame_parent"])
    d.n_dofs = len(pack["joints_lower"])
    d.n_links = len(pack["mesh_frame"])
    d.n_cand = len(pack["cand"])
    d.n_spheres = len(pack["spheres"])
    d.frame_parent = arr("frame_parent", np.int32)
    d.frame_jtype = arr("frame_jtype", np.int32)
    d.frame_dof = arr("frame_dof", np.int32)
    d.frame_axis = arr("frame_axis", np.float32)
    d.frame_joff = arr("frame_joff", np.float32)
    d.frame_loff = arr("frame_loff", np.float32)
    d.link_frame = arr("mesh_frame", np.int32)
    d.face_offset = arr("face_offset", np.int32)
    d.face_verts = arr("face_verts", np.float32)
    d.cand = arr("cand", np.float32)
    d.cand_normal = arr("cand_normal", np.float32)
    d.cand_link = arr("cand_link", np.int32)
    d.spheres = arr("spheres", np.float32)
    d.sphere_link = arr("sphere_link", np.int32)
    d.joints_lower = arr("joints_lower", np.float32)
    d.joints_upper = arr("joints_upper", np.float32)
    if "mlp_w1" in pack:  # joint-coupling network (Robotiq 2F-140)
        d.mlp_hidden, d.mlp_out = len(pack["mlp_w1"]), len(pack["mlp_b2"])
        d.mlp_lo, d.mlp_hi = float(pack["mlp_clamp"][0]), float(pack["mlp_clamp"][1])
        d.mlp_w1, d.mlp_b1 = arr("mlp_w1", np.float32), arr("mlp_b1", np.float32)
        d.mlp_w2, d.mlp_b2 = arr("mlp_w2", np.float32), arr("mlp_b2", np.float32)
    if "frame_gain" in pack:  # mimic joints (ability hand)
        d.frame_gain = arr("frame_gain", np.float32)
    return d, keep


def create_hand(cdll, pack):
    d, keep = make_desc(pack)
    h = ctypes.c_void_p()
    _lib.check(cdll.gqp_hand_create(ctypes.byref(d), ctypes.byref(h)), cdll)
    return h
