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


def joint_vector(pack, theta):
    """utils/fk.py:23-51 for a hand with a coupling network: theta (B,1) -> q (B, 1 + mlp_out) = [c, W2 tanh(w1 c + b1) + b2],
    c = clamp(theta); host numpy (the kernels evaluate the same inside the FK)."""
    th = np.asarray(theta, np.float32).reshape(-1, 1)
    c = np.clip(th, pack["mlp_clamp"][0], pack["mlp_clamp"][1])
    h = np.tanh(c * np.asarray(pack["mlp_w1"], np.float32).reshape(1, -1) + np.asarray(pack["mlp_b1"], np.float32))
    return np.concatenate([c, h @ np.asarray(pack["mlp_w2"], np.float32).T + np.asarray(pack["mlp_b2"], np.float32)], 1)


def expanded_pack(pack):
    """The same hand with the EXPANDED joint vector q as its parameters and no coupling network: what the reference's
    per-hand jacobian_fnc differentiates (hands/robotiq2.py:17-20 takes pk's Jacobian at the expanded joints)."""
    n_q = 1 + len(pack["mlp_b2"])
    out = {k: v for k, v in pack.items() if not k.startswith("mlp_")}
    out["joints_lower"] = np.full(n_q, -1e30, np.float32)
    out["joints_upper"] = np.full(n_q, 1e30, np.float32)
    out["default_state"] = np.zeros(n_q, np.float32)
    names = [""] * n_q
    for f, d in enumerate(pack["frame_dof"]):
        if int(pack["frame_jtype"][f]) != 0 and d >= 0:
            names[int(d)] = str(pack["frame_joint"][f])
    out["joint_names"] = np.array(names)
    return out


def reference_jacobian_columns(pack):
    """For a hand with a coupling network the reference does NOT differentiate through the network: its jacobian_fnc
    (hands/robotiq2.py:17-20) returns column 0 minus column 3 of pk's Jacobian at the expanded joints, columns counted in the
    chain's joint order.  -> [(column of q, weight), ...]"""
    movable = [f for f in range(len(pack["frame_jtype"])) if int(pack["frame_jtype"][f]) != 0]
    return [(int(pack["frame_dof"][movable[0]]), 1.0), (int(pack["frame_dof"][movable[3]]), -1.0)]


def surface_points(pack, n_surface_points=512, seed=42):
    """The hand surface samples E_wall reads (HandModel.get_surface_points, hand_model.py:1042-1071): per link
    int(area / total_area * n) points, the remainder to the first link, each link's set = farthest point sampling (start
    index 0) of 100x as many area-uniform samples (hand_model.py:603-629).  The reference draws the dense cloud with
    pytorch3d under torch.manual_seed(42); here numpy's RandomState(seed) draws it, so the point SET differs from the
    reference's while its construction is the same.  Setup-time host work.  -> (points (n,3) float32 in link frames,
    link (n) int32 ascending)."""
    fv = np.asarray(pack["face_verts"], np.float64).reshape(-1, 3, 3)
    off = np.asarray(pack["face_offset"])
    L = len(off) - 1
    tri_area = 0.5 * np.linalg.norm(np.cross(fv[:, 1] - fv[:, 0], fv[:, 2] - fv[:, 0]), axis=1)
    areas = np.array([tri_area[off[l]:off[l + 1]].sum() for l in range(L)])
    num = [int(a / areas.sum() * n_surface_points) for a in areas]
    num[0] += n_surface_points - sum(num)
    pts, link = [], []
    for l in range(L):
        k = num[l]
        if k == 0:
            continue
        rng = np.random.RandomState(seed)
        a = tri_area[off[l]:off[l + 1]]
        f = rng.choice(len(a), size=100 * k, p=a / a.sum())
        u, v = rng.rand(100 * k), rng.rand(100 * k)
        s = np.sqrt(u)
        t = fv[off[l]:off[l + 1]][f]
        dense = ((1 - s)[:, None] * t[:, 0] + (s * (1 - v))[:, None] * t[:, 1] + (s * v)[:, None] * t[:, 2]).astype(np.float32)
        sel, mind, cur = np.zeros(k, np.int64), np.full(len(dense), np.inf, np.float32), 0
        for i in range(k):
            sel[i] = cur
            d = dense - dense[cur]
            mind = np.minimum(mind, (d * d).sum(1))
            cur = int(np.argmax(mind))
        pts.append(dense[sel])
        link.append(np.full(k, l, np.int32))
    return np.concatenate(pts, 0).astype(np.float32), np.concatenate(link, 0)
