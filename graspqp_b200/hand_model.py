"""HandModel — the reference's class (graspqp/src/graspqp/core/hand_model.py:395-1267) for the hot path, over the
sm_100a kernels.  Same attribute and method names, so scripts/fit.py's loop, MalaStar and calculate_energy run unchanged:

    set_parameters(hand_pose, contact_point_indices)   hand_model.py:833-873
    cal_distance(x)                                     :875-987   (SDF_BACKEND TORCHSDF semantics)
    self_penetration()                                  :989-1040
    fk(joint_angles)                                    :762-766   (returns world link transforms here)
    attributes: n_dofs, n_contact_candidates, joints_lower/upper, hand_pose, contact_points, contact_normals,
                contact_point_indices, global_rotation, global_translation, current_status, device

What is NOT here (out of the hot path's scope): URDF/mesh parsing at construction (done once by
tools/build_hand_pack.py into assets/<hand>.npz), plotting, retargeting helpers.
"""
import numpy as np
import torch

from . import _lib, hand_pack, ops


class HandModel:
    def __init__(self, hand="allegro", device="cuda", pack=None):
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("graspqp_b200.HandModel runs on CUDA devices only (no CPU fallback)")
        self.pack = pack if pack is not None else hand_pack.load_pack(hand)
        self.name = str(self.pack["name"])
        with torch.cuda.device(self.device):
            self.handle = hand_pack.create_hand(_lib.lib(), self.pack)
        t = lambda k, dt=torch.float32: torch.as_tensor(np.asarray(self.pack[k]), dtype=dt, device=self.device)
        self.n_dofs = len(self.pack["joints_lower"])
        self.n_links = len(self.pack["mesh_frame"])
        self.n_contact_candidates = len(self.pack["cand"])
        self.joints_lower, self.joints_upper = t("joints_lower"), t("joints_upper")
        self.default_state = t("default_state")
        self.contact_candidates, self.normal_candidates = t("cand"), t("cand_normal")
        self.global_index_to_link_index = t("cand_link", torch.long)
        self.penetration_keypoints = t("spheres")
        self._actuated_joints_names = [str(s) for s in self.pack["joint_names"]]
        self._contact_links = [str(s) for s in self.pack["mesh_names"]]
        self.mesh_names = self._contact_links
        ax = {"x": (1, 0, 0), "y": (0, 1, 0), "z": (0, 0, 1), "-x": (-1, 0, 0), "-y": (0, -1, 0), "-z": (0, 0, -1)}
        axes = [str(a) for a in self.pack["axes"]]
        vec = lambda a: torch.tensor(ax[a], dtype=torch.float32, device=self.device)
        self.forward_axis, self.up_axis, self.grasp_axis = vec(axes[0]), vec(axes[1]), vec(axes[2])  # hand_model.py:459-470
        self._surf = None  # hand surface samples for get_surface_points (E_wall), built on first use
        self.hand_pose = None
        self.contact_point_indices = None
        self.contact_points = self.contact_normals = None
        self.global_translation = self.global_rotation = None
        self.current_status = None  # (Bt, n_links, 12) world transforms [R|t] of the mesh links

    def __del__(self):
        try:
            if getattr(self, "handle", None):
                _lib.lib().gqp_hand_destroy(self.handle)
                self.handle = None
        except Exception:
            pass

    # ---- kinematics --------------------------------------------------------------------------------------------
    def fk(self, joint_angles=None):
        """hand_model.py:762-766.  Returns the WORLD transforms of the mesh links for the current global pose
        (the reference returns hand-frame Transform3d objects; its consumers compose the global pose afterwards)."""
        pose = self.hand_pose if joint_angles is None else torch.cat([self.hand_pose[:, :9], joint_angles], 1)
        return ops.FkFn.apply(pose, self)

    def set_parameters(self, hand_pose, contact_point_indices=None, env_mask=None):
        if env_mask is not None:  # hand_model.py:846-850
            with torch.no_grad():
                hp = torch.where(env_mask.unsqueeze(-1), hand_pose, self.hand_pose)
            self.hand_pose = hp.detach().requires_grad_(True)
        else:
            self.hand_pose = hand_pose.clone()
        if self.hand_pose.requires_grad:
            self.hand_pose.retain_grad()
        if torch.isnan(self.hand_pose).any():
            raise ValueError("nan in hand_pose")
        self.global_translation = self.hand_pose[:, 0:3]
        self.current_status = ops.FkFn.apply(self.hand_pose, self)
        # an attribute for callers (E_prior reads it, energy.py:68-74); the kernels carry R_g inside current_status
        x, y = self.hand_pose[:, 3:6], self.hand_pose[:, 6:9]
        x = x / x.norm(dim=-1, keepdim=True)
        y = y - (x * y).sum(-1, keepdim=True) * x
        y = y / y.norm(dim=-1, keepdim=True)
        self.global_rotation = torch.stack((x, y, torch.cross(x, y, dim=-1)), dim=-1)
        gather_idx = self.contact_point_indices
        if contact_point_indices is not None:
            # hand_model.py:816-831: with env_mask the MEMBER is merged, but the contact points of this call are gathered
            # with the ARGUMENT (the fresh indices of every row) — kept as the reference does it
            gather_idx = contact_point_indices
            if env_mask is not None and self.contact_point_indices is not None:
                self.contact_point_indices = torch.where(env_mask.unsqueeze(-1), contact_point_indices,
                                                         self.contact_point_indices)
            else:
                self.contact_point_indices = contact_point_indices.clone()
        if gather_idx is not None:
            self.contact_points, self.contact_normals = ops.ContactsFn.apply(self.current_status, gather_idx, self)

    # ---- distances ---------------------------------------------------------------------------------------------
    def cal_distance(self, x):
        """hand_model.py:875-987: signed depth of world points x (Bt,P,3) w.r.t. the hand, max over links, positive
        inside.  Differentiable w.r.t. hand_pose."""
        return ops.HandDistanceFn.apply(self.current_status, x, self)

    def self_penetration(self):
        """hand_model.py:989-1040."""
        return ops.SelfPenFn.apply(self.current_status, self.hand_pose, self)

    def get_surface_points(self):
        """hand_model.py:1042-1071: world positions of the hand surface samples (Bt, n, 3); the sample table is
        hand_pack.surface_points (512 samples, built the reference's way from a different random stream)."""
        if self._surf is None:
            pts, link = hand_pack.surface_points(self.pack)
            self._surf = (torch.as_tensor(pts, device=self.device), torch.as_tensor(link, device=self.device).long())
        pts, link = self._surf
        tf = self.current_status.view(self.current_status.shape[0], self.n_links, 3, 4)[:, link]  # (Bt, n, 3, 4)
        return (tf[..., :3] @ pts.unsqueeze(-1)).squeeze(-1) + tf[..., 3]

    def get_contact_candidates(self):
        """hand_model.py:1220-1267: world positions of ALL contact candidates (Bt, C, 3)."""
        B = self.hand_pose.shape[0]
        idx = torch.arange(self.n_contact_candidates, device=self.device, dtype=torch.int32).repeat(B, 1)
        return ops.ContactsFn.apply(self.current_status, idx, self)[0]


# graspqp/src/graspqp/hands/__init__.py:13-28 (registry names -> hand packs).  "schunk2" is not built: the reference ships no
# contact_infos.json for it (HandModel would sample it with trimesh at construction) and its meshes are COLLADA files.
_REGISTRY = {"robotiq3": "robotiq3", "panda": "panda", "allegro": "allegro", "ability_hand": "ability_hand",
             "shadow_hand": "shadow_hand", "robotiq2": "robotiq2"}
AVAILABLE_HANDS = list(_REGISTRY.keys())


def get_hand_model(hand_name, device, **kwargs):
    if hand_name not in _REGISTRY:
        raise KeyError(f"hand {hand_name!r}: available {AVAILABLE_HANDS}")
    if kwargs.get("grasp_type") not in (None, "all", "default"):
        raise NotImplementedError("grasp_type filters of the contact links (eigengrasps.json) are not compiled into the packs")
    return HandModel(_REGISTRY[hand_name], device)
