"""TEST INFRASTRUCTURE — NOT PRODUCT CODE.

PyTorch (CPU, autograd) restatement of the reference HandModel methods on the hot path, op by op, so that autograd
produces the same gradient graph the reference's `new_energy.sum().backward()` walks (scripts/fit.py:440).

Follows
* graspqp/src/graspqp/core/hand_model.py:762-766 (fk), :787-873 (set_parameters/_set_contact_idxs),
  :875-987 (cal_distance, TORCHSDF branch), :989-1040 (self_penetration), :1220-1267 (get_contact_candidates)
* thirdparty/pytorch_kinematics/pytorch_kinematics/chain.py:146-168, frame.py:98-114 (incl. the +-0.01 rad
  perturbation), transforms/rotation_conversions.py:43-71, transforms/transform3d.py:258-272, 369-444
* graspqp/src/graspqp/utils/transforms.py:5-13 -> roma.special_gramschmidt.  roma==1.5.4 is NOT vendored in the
  reference and not installed here: restated from its published algorithm (classic Gram-Schmidt of the two columns,
  z = x cross y, R = [x y z] as columns).  PARITY UNPINNED for this function.

The hand description comes from graspqp_b200/assets/<hand>.npz (tools/build_hand_pack.py).
"""
import numpy as np
import torch

from .sdf import compute_sdf_port


class _SdfFn(torch.autograd.Function):
    """thirdparty/TorchSDF/torchsdf/sdf.py:42-68: only dist^2 is differentiable, backward = 2 g (p - c)."""

    @staticmethod
    def forward(ctx, points, face_verts):
        d2, sg, nr, cl = compute_sdf_port(points.detach().to(torch.float32).numpy(), face_verts)
        dt = points.dtype
        d2, sg, nr, cl = torch.from_numpy(d2).to(dt), torch.from_numpy(sg), torch.from_numpy(nr).to(dt), torch.from_numpy(cl).to(dt)
        if dt == torch.float64:  # fp64 truth runs: recompute dist^2 from the fp32 winner's closest point in fp64
            d2 = ((points.detach() - cl) ** 2).sum(-1)
        ctx.save_for_backward(points.detach(), cl)
        ctx.mark_non_differentiable(sg, nr, cl)
        return d2, sg, nr, cl

    @staticmethod
    def backward(ctx, g, *_):
        p, c = ctx.saved_tensors
        return 2.0 * g.unsqueeze(-1) * (p - c), None


def compute_sdf(points, face_verts):
    return _SdfFn.apply(points, face_verts)


def special_gramschmidt(M):
    x, y = M[..., 0], M[..., 1]
    x = x / torch.clamp_min(torch.norm(x, dim=-1, keepdim=True), 0.0)
    y = y - torch.sum(x * y, dim=-1, keepdim=True) * x
    y = y / torch.clamp_min(torch.norm(y, dim=-1, keepdim=True), 0.0)
    z = torch.cross(x, y, dim=-1)
    return torch.stack((x, y, z), dim=-1)


def quaternion_to_matrix(q):
    r, i, j, k = torch.unbind(q, -1)
    two_s = 2.0 / (q * q).sum(-1)
    o = torch.stack((1 - two_s * (j * j + k * k), two_s * (i * j - k * r), two_s * (i * k + j * r),
                     two_s * (i * j + k * r), 1 - two_s * (i * i + k * k), two_s * (j * k - i * r),
                     two_s * (i * k - j * r), two_s * (j * k + i * r), 1 - two_s * (i * i + j * j)), -1)
    return o.reshape(q.shape[:-1] + (3, 3))


def _mat44(R=None, t=None, B=1, dtype=torch.float32, device="cpu"):
    m = torch.eye(4, dtype=dtype, device=device).repeat(B, 1, 1)
    if R is not None:
        m = m.clone()
        m[:, :3, :3] = R
    if t is not None:
        m[:, :3, 3] = t
    return m


class HandOracle:
    def __init__(self, pack, dtype=torch.float32, device="cpu", sdf=None):
        """device / sdf: the GPU baseline leg (baseline/gpu_reference.py) runs this same restatement on CUDA tensors with
        sdf = the reference's own torchsdf.compute_sdf; the default is the CPU port of that kernel."""
        self.dtype, self.device = dtype, torch.device(device)
        self.sdf = sdf if sdf is not None else compute_sdf
        g = lambda k: torch.from_numpy(np.asarray(pack[k])).to(self.device)
        self.parent = [int(v) for v in pack["frame_parent"]]
        self.jtype = [int(v) for v in pack["frame_jtype"]]
        self.frame_dof = [int(v) for v in pack["frame_dof"]]
        # mimic joints (hands/ability_hand.py:10-29): joint value = gain * parameter
        self.frame_gain = [float(v) for v in pack["frame_gain"]] if "frame_gain" in pack else [1.0] * len(self.parent)
        self.axis = g("frame_axis").to(dtype)
        to44 = lambda a: torch.cat([a, torch.tensor([0, 0, 0, 1.0], device=self.device).expand(a.shape[0], 1, 4)], 1).to(dtype)
        self.joff, self.loff = to44(g("frame_joff")), to44(g("frame_loff"))
        self.mesh_frame = [int(v) for v in pack["mesh_frame"]]
        self.face_verts = np.asarray(pack["face_verts"], np.float32)
        self.face_offset = [int(v) for v in pack["face_offset"]]
        if sdf is not None:  # link meshes as device tensors for torchsdf
            self.face_verts_dev = [torch.from_numpy(self.face_verts[a:b].reshape(-1, 3, 3)).to(self.device).contiguous()
                                   for a, b in zip(self.face_offset[:-1], self.face_offset[1:])]
        self.cand, self.cand_normal = g("cand").to(dtype), g("cand_normal").to(dtype)
        self.cand_link = g("cand_link").long()
        self.spheres, self.sphere_link = g("spheres").to(dtype), g("sphere_link").long()
        self.joints_lower, self.joints_upper = g("joints_lower").to(dtype), g("joints_upper").to(dtype)
        self.n_dofs = len(self.joints_lower)
        self.n_contact_candidates = len(self.cand)
        self.L = len(self.mesh_frame)
        self.hand_pose = None
        # HandModel.grasp_axis (hand_model.py:459-470) and the surface samples of get_surface_points (:1042-1071); the
        # samples are injected (graspqp_b200.hand_pack.surface_points builds a table the reference's way)
        ax = {"x": (1, 0, 0), "y": (0, 1, 0), "z": (0, 0, 1), "-x": (-1, 0, 0), "-y": (0, -1, 0), "-z": (0, 0, -1)}
        self.grasp_axis = torch.tensor(ax[str(pack["axes"][2])], dtype=dtype, device=self.device)
        self.surface_pts = self.surface_link = None
        # hands/robotiq2.py:9-13 + utils/fk.py:23-51: Linear(1,10)-Tanh-Linear(10,5) joint-coupling network
        self.mlp = None
        if "mlp_w1" in pack:
            self.mlp = (g("mlp_w1").to(dtype).reshape(-1, 1), g("mlp_b1").to(dtype), g("mlp_w2").to(dtype),
                        g("mlp_b2").to(dtype), float(pack["mlp_clamp"][0]), float(pack["mlp_clamp"][1]))

    def joint_vector(self, theta):
        """utils/fk.py:32-39: cat([clamp(theta), net(clamp(theta))]); frame_dof indexes these columns."""
        if self.mlp is None:
            return theta
        w1, b1, w2, b2, lo, hi = self.mlp
        c = torch.clamp(theta, min=lo, max=hi)
        return torch.cat([c, torch.tanh(c @ w1.T + b1) @ w2.T + b2], -1)

    # pk/chain.py:146-168 + frame.py:98-114
    def fk(self, theta):
        B = theta.shape[0]
        theta = self.joint_vector(theta)
        trans, out = [None] * len(self.parent), [None] * len(self.parent)
        for f, par in enumerate(self.parent):
            world = torch.eye(4, dtype=self.dtype, device=self.device).expand(B, 4, 4) if par < 0 else trans[par]
            d = self.frame_dof[f]
            if self.jtype[f] == 1 and d >= 0:
                th = theta[:, d:d + 1] * self.frame_gain[f]
                th = th + 2 * ((th > 0).to(self.dtype) - 0.5) * 1e-2
                quat = torch.cat([torch.cos(th / 2), torch.sin(th / 2) * self.axis[f]], -1)
                jt = _mat44(quaternion_to_matrix(quat), None, B, self.dtype, self.device)
            elif self.jtype[f] == 2 and d >= 0:
                jt = _mat44(None, theta[:, d:d + 1] * self.frame_gain[f] * self.axis[f], B, self.dtype, self.device)
            elif self.jtype[f] == 1:  # movable joint that is not in the parameter vector: th_dict.get(name, 0.0)
                th = torch.zeros(B, 1, dtype=self.dtype, device=self.device)
                th = th + 2 * ((th > 0).to(self.dtype) - 0.5) * 1e-2
                quat = torch.cat([torch.cos(th / 2), torch.sin(th / 2) * self.axis[f]], -1)
                jt = _mat44(quaternion_to_matrix(quat), None, B, self.dtype, self.device)
            else:
                jt = torch.eye(4, dtype=self.dtype, device=self.device).expand(B, 4, 4)
            trans[f] = world @ (self.joff[f] @ jt)
            out[f] = trans[f] @ self.loff[f]
        return out

    def link_matrix(self, li):
        return self.current_status[self.mesh_frame[li]]

    def set_parameters(self, hand_pose, contact_point_indices, env_mask=None):
        if env_mask is not None:  # hand_model.py:845-850
            with torch.no_grad():
                self.hand_pose = torch.where(env_mask.unsqueeze(-1), hand_pose, self.hand_pose)
            self.hand_pose.requires_grad = True
            self.hand_pose.retain_grad()
        else:
            self.hand_pose = hand_pose.clone()
        if self.hand_pose.requires_grad:
            self.hand_pose.retain_grad()
        self.global_translation = self.hand_pose[:, 0:3]
        self.global_rotation = special_gramschmidt(torch.stack([self.hand_pose[:, 3:6], self.hand_pose[:, 6:9]], -1))
        self.current_status = self.fk(self.hand_pose[:, 9:])
        if env_mask is None:  # hand_model.py:816-822
            self.contact_point_indices = contact_point_indices.clone()
        else:
            self.contact_point_indices = torch.where(env_mask.unsqueeze(-1), contact_point_indices,
                                                     self.contact_point_indices)
        pts, nrm = [], []
        for li in range(self.L):
            m = self.cand_link == li
            if m.sum() == 0:
                continue
            T = self.link_matrix(li)
            R, t = T[:, :3, :3], T[:, :3, 3]
            pts.append(self.cand[m] @ R.transpose(1, 2) + t.unsqueeze(1))  # transform_points
            nrm.append(self.cand_normal[m] @ torch.linalg.inv(R))  # transform_normals: n @ R^-1
        pts = torch.cat(pts, -2) @ self.global_rotation.transpose(1, 2) + self.global_translation.unsqueeze(1)
        nrm = torch.cat(nrm, -2) @ self.global_rotation.transpose(1, 2)
        # hand_model.py:830-831: the gather uses the ARGUMENT (with env_mask: the fresh indices of every row), not the
        # merged member
        idx = contact_point_indices.unsqueeze(-1).expand(-1, -1, 3)
        self.contact_points = pts.gather(1, idx)
        self.contact_normals = nrm.gather(1, idx)

    def set_surface_points(self, pts, link):
        self.surface_pts = torch.as_tensor(pts).to(self.dtype).to(self.device)
        self.surface_link = torch.as_tensor(link).long().to(self.device)

    def get_surface_points(self):
        """hand_model.py:1042-1071: link-frame samples -> world, links in mesh order"""
        out = []
        for li in range(self.L):
            m = self.surface_link == li
            if m.sum() == 0:
                continue
            T = self.link_matrix(li)
            out.append(self.surface_pts[m] @ T[:, :3, :3].transpose(1, 2) + T[:, :3, 3].unsqueeze(1))
        pts = torch.cat(out, -2)
        return pts @ self.global_rotation.transpose(1, 2) + self.global_translation.unsqueeze(1)

    def cal_distance(self, x, per_link=False):
        x = (x - self.global_translation.unsqueeze(1)) @ self.global_rotation
        dis = []
        for li in range(self.L):
            fv = self.face_verts[self.face_offset[li]:self.face_offset[li + 1]]
            if len(fv) == 0:
                continue
            if self.sdf is not compute_sdf:
                fv = self.face_verts_dev[li]
            T = self.link_matrix(li)
            xl = ((x - T[:, :3, 3].unsqueeze(1)) @ T[:, :3, :3]).reshape(-1, 3)
            d2, sg, _, _ = self.sdf(xl, fv)
            d = torch.sqrt(d2 + 1e-8) * (-sg)
            dis.append(d.reshape(x.shape[0], x.shape[1]))
        dis = torch.stack(dis, 0)
        return dis if per_link else torch.max(dis, dim=0)[0]

    def self_penetration(self):
        B = self.global_translation.shape[0]
        if len(self.spheres) == 0:
            return torch.zeros(B, dtype=self.dtype, device=self.device)
        points, radii = [], []
        for li in range(self.L):
            m = self.sphere_link == li
            if m.sum() == 0:
                continue
            T = self.link_matrix(li)
            p = self.spheres[m, :3] @ T[:, :3, :3].transpose(1, 2) + T[:, :3, 3].unsqueeze(1)
            points.append(p @ self.global_rotation.transpose(1, 2) + self.global_translation.unsqueeze(1))
            radii.append(self.spheres[m, 3])
        all_pts, all_r = torch.cat(points, -2), torch.cat(radii)
        counter = 0
        distances = torch.zeros(B, len(points) - 1, dtype=self.dtype, device=self.device)
        for idx, ls in enumerate(points[:-1]):
            r = all_r[counter:counter + ls.shape[1]]
            counter += ls.shape[1]
            other, other_r = all_pts[:, counter:], all_r[counter:]
            dis = (ls.unsqueeze(1) - other.unsqueeze(2) + 1e-13).norm(dim=-1)
            pen = dis - (r.view(1, -1) + other_r.view(-1, 1))
            distances[:, idx] = pen.min(1)[0].min(1)[0]
        return -distances.clamp(max=0).sum(1)
