"""(f2) Export path — scripts/fit.py:224-300 (`export_poses`) with HandModel.get_req_joint_velocities
(core/hand_model.py:1155-1218: contact Jacobian + damped pseudo-inverse, hand_model.py:46-54) and the `.dexgrasp.pt` dict
(fit.py:203-221, 288-299) that graspqp_isaaclab/utils/data.py:105-171 and scripts/vis/visualize_result.py read.

The reference assembles the contact Jacobian from pytorch_kinematics' vectorised link Jacobian (pk/chain.py:197-252) plus
omega x r per contact, in the hand frame, and rotates the desired contact motion into that frame.  The linear part of that
Jacobian IS d(contact point)/d(joint parameter) (checked against pk in tests/test_export.py), and with the orthogonal
global rotation R_g  pinv_l(J_hand) R_g' v == pinv_l(R_g J_hand) v, so here the WORLD-frame Jacobian comes straight from
the kernels that already carry the analytic J' of the kinematic chain: row (contact c, axis k) of J is gqp_fk_backward
applied to a one-hot cotangent on contact c — 3 n_c passes of two launches each, once per export, nothing new on the
per-iteration path.  Fixed-gain couplings (Panda's mirrored fingers, the ability hand's mimic joints) are inside those kernels,
which is what the reference's per-hand `jacobian_fnc` (hands/panda.py:17-26, hands/ability_hand.py:36-42) sums by hand; for the
Robotiq 2F-140 the reference's jacobian_fnc is NOT the derivative through its coupling network (hands/robotiq2.py:17-20) and
`contact_jacobian` reproduces what it returns instead.
"""
import ctypes
import os

import numpy as np

from . import _lib


def pinv(A, lam=1e-3):
    """core/hand_model.py:46-54, torch"""
    import torch
    m, n = A.shape[-2:]
    if m < n:
        return A.mT @ torch.linalg.inv(A @ A.mT + lam * torch.eye(m, device=A.device, dtype=A.dtype))
    return torch.linalg.inv(A.mT @ A + lam * torch.eye(n, device=A.device, dtype=A.dtype)) @ A.mT


def rotmat_to_quat_wxyz(R):
    """(B,3,3) torch -> unit quaternions (w, x, y, z), w >= 0 (fit.py:261-263 takes roma.rotmat_to_unitquat's xyzw and
    reorders; q and -q are the same rotation).  Shepperd's method: divide by the largest component."""
    import torch
    m00, m11, m22 = R[:, 0, 0], R[:, 1, 1], R[:, 2, 2]
    q = 0.5 * torch.sqrt(torch.clamp(torch.stack([1 + m00 + m11 + m22, 1 + m00 - m11 - m22, 1 - m00 + m11 - m22,
                                                  1 - m00 - m11 + m22], 1), min=0))  # |w|, |x|, |y|, |z|
    a, b, c = R[:, 2, 1] - R[:, 1, 2], R[:, 0, 2] - R[:, 2, 0], R[:, 1, 0] - R[:, 0, 1]
    d, e, f = R[:, 0, 1] + R[:, 1, 0], R[:, 0, 2] + R[:, 2, 0], R[:, 1, 2] + R[:, 2, 1]
    cand = torch.stack([torch.stack([4 * q[:, 0] ** 2, a, b, c], 1), torch.stack([a, 4 * q[:, 1] ** 2, d, e], 1),
                        torch.stack([b, d, 4 * q[:, 2] ** 2, f], 1), torch.stack([c, e, f, 4 * q[:, 3] ** 2], 1)], 1)
    k = q.argmax(1)
    rows = torch.arange(len(R), device=R.device)
    out = cand[rows, k] / (4 * q[rows, k]).unsqueeze(1)
    out = torch.where(out[:, :1] < 0, -out, out)
    return out / out.norm(dim=1, keepdim=True)


class Exporter:
    """export_poses over a FitEngine's device state (TorchBackend)."""

    def __init__(self, eng):
        import torch
        self.eng = eng
        self.torch = torch

    def _t(self, a):
        """backend array -> torch tensor (the emulator backend of the tests holds numpy arrays)"""
        return a if self.torch.is_tensor(a) else self.torch.as_tensor(np.asarray(a))

    # ---- world-frame contact Jacobian d(contact point)/d(joint parameter): (Bt, n, 3, dof) ----
    def contact_jacobian(self, idx):
        """(Bt, n, 3, dof).  Hands whose parameters reach the joints through fixed gains (Panda, ability hand) or directly:
        the chain-rule derivative, which is what their jacobian_fnc sums by hand (hands/panda.py:17-26,
        hands/ability_hand.py:36-42).  The Robotiq 2F-140 (coupling network): the reference's jacobian_fnc does not
        differentiate the network, it returns pk's column of `finger_joint` minus that of `right_outer_knuckle_joint` at the
        expanded joints (hands/robotiq2.py:17-20) — reproduced as such."""
        eng = self.eng
        if "mlp_w1" not in eng.pack:
            return self._jacobian_rows(eng.hand, eng.pose, eng.dof, idx)
        from . import hand_pack
        be = eng.be
        if getattr(self, "_qhand", None) is None:
            self._qpack = hand_pack.expanded_pack(eng.pack)
            self._qhand = hand_pack.create_hand(eng.lib, self._qpack)
        pose = be.to_host(eng.pose)
        q = hand_pack.joint_vector(eng.pack, pose[:, 9:])
        Jq = self._jacobian_rows(self._qhand, be.from_host(np.concatenate([pose[:, :9], q], 1), np.float32), q.shape[1], idx)
        J = None
        for col, wgt in hand_pack.reference_jacobian_columns(eng.pack):
            J = wgt * Jq[..., col:col + 1] if J is None else J + wgt * Jq[..., col:col + 1]
        return J

    def close(self):
        if getattr(self, "_qhand", None) is not None:
            self.eng.lib.gqp_hand_destroy(self._qhand)
            self._qhand = None

    def _jacobian_rows(self, hand, pose, dof, idx):
        """row (contact c, axis k) of d(contact point)/d(joint parameter) = the FK backward kernels on a one-hot cotangent"""
        eng, lib, be = self.eng, self.eng.lib, self.eng.be
        p, st = be.ptr, be.stream
        Bt, n = idx.shape
        link_tf = be.empty((Bt, eng.L, 12), np.float32)
        _lib.check(lib.gqp_fk_forward(hand, p(pose), Bt, p(link_tf), st), lib)
        g_cp, g_cn = be.empty((Bt, n, 3), np.float32), be.empty((Bt, n, 3), np.float32)
        g_tf, g_pose = be.empty((Bt, eng.L, 12), np.float32), be.empty((Bt, 9 + dof), np.float32)
        J = be.empty((Bt, n, 3, dof), np.float32)
        for c in range(n):
            for k in range(3):
                g_cp[...] = 0
                g_cp[:, c, k] = 1.0
                g_tf[...] = 0
                _lib.check(lib.gqp_contacts_backward(hand, p(idx), p(g_cp), p(g_cn), Bt, n, p(g_tf), st), lib)
                _lib.check(lib.gqp_fk_backward(hand, p(pose), p(g_tf), None, Bt, p(g_pose), st), lib)
                J[:, c, k, :] = g_pose[:, 9:]
        return J

    def contacts_and_object(self, idx):
        """contact points / object distance / outward object normal at the CURRENT state for the contact table idx"""
        eng, lib, be = self.eng, self.eng.lib, self.eng.be
        p, st = be.ptr, be.stream
        Bt, n = idx.shape
        link_tf = be.empty((Bt, eng.L, 12), np.float32)
        _lib.check(lib.gqp_fk_forward(eng.hand, p(eng.pose), Bt, p(link_tf), st), lib)
        cp, cn = be.empty((Bt, n, 3), np.float32), be.empty((Bt, n, 3), np.float32)
        _lib.check(lib.gqp_contacts_forward(eng.hand, p(link_tf), p(idx), Bt, n, p(cp), p(cn), st), lib)
        dist, nrm = be.empty((Bt, n), np.float32), be.empty((Bt, n, 3), np.float32)
        clst, sign = be.empty((Bt, n, 3), np.float32), be.empty((Bt, n), np.int32)
        _lib.check(lib.gqp_object_distance_forward(eng.objset, eng.batch_each, n, p(cp), 1.0, p(dist), p(nrm), p(clst),
                                                   p(sign), st), lib)
        return cp, dist, nrm

    def get_req_joint_velocities(self, moving_directions, idx, return_ee_vel=False):
        """hand_model.py:1155-1218 (coupled=True): theta = pinv(J) v, residuals, ee velocities; world frame throughout"""
        J = self._t(self.contact_jacobian(idx)).flatten(1, 2)  # (Bt, 3n, dof)
        v = self._t(moving_directions).flatten(1, 2).unsqueeze(-1)
        theta = pinv(J) @ v
        ee = J @ theta
        res = ((ee - v) ** 2).squeeze(-1)
        if return_ee_vel:
            return theta.squeeze(-1), res, ee.view(moving_directions.shape)
        return theta.squeeze(-1), res

    def grasp_data(self, grasp_type=None):
        """-> list of dicts, one per object (fit.py:253-299), tensors on the CPU"""
        eng, torch = self.eng, self.torch
        pack, B = eng.pack, eng.batch_each
        idx = eng.idx
        tt = self._t
        _, d, nrm = [tt(a) for a in self.contacts_and_object(idx)]
        dtheta, _, _ = self.get_req_joint_velocities(-(5 * (nrm * d.unsqueeze(-1).abs())), idx, True)       # fit.py:232-236
        all_idx = eng.be.from_host(np.tile(np.arange(eng.n_cand, dtype=np.int32), (eng.Bt, 1)), np.int32)
        _, d_f, nrm_f = [tt(a) for a in self.contacts_and_object(all_idx)]                                                    # :238-243
        dtheta_full, _ = self.get_req_joint_velocities(-(5 * (nrm_f * d_f.unsqueeze(-1).abs())), all_idx)
        dtheta_off, _, _ = self.get_req_joint_velocities(-(5 * nrm * (d.unsqueeze(-1).abs() + 0.005)), idx, True)  # :247-252
        pose, energy = tt(eng.pose).detach().cpu(), tt(eng.energy).detach().cpu()
        x, y = pose[:, 3:6], pose[:, 6:9]  # utils/transforms.py:5-13 (special Gram-Schmidt, columns)
        x = x / x.norm(dim=-1, keepdim=True)
        y = y - (x * y).sum(-1, keepdim=True) * x
        y = y / y.norm(dim=-1, keepdim=True)
        quat = rotmat_to_quat_wxyz(torch.stack((x, y, torch.cross(x, y, dim=-1)), dim=-1))
        names = [str(s) for s in pack["joint_names"]]
        out = []
        for o in range(eng.n_obj):
            s = slice(o * B, (o + 1) * B)
            params = {names[j]: pose[s, 9 + j].clone() for j in range(eng.dof)}
            params["root_pose"] = torch.cat([pose[s, :3], quat[s]], 1)
            vel = lambda t: {names[j]: t[s, j].detach().cpu().clone() for j in range(eng.dof)}
            out.append({"values": energy[s].clone(), "parameters": params, "grasp_velocities": vel(dtheta),
                        "full_grasp_velocities": vel(dtheta_full), "grasp_velocities_off": vel(dtheta_off),
                        "contact_idx": tt(idx)[s].detach().cpu().long(), "grasp_type": grasp_type,
                        "contact_links": [str(n) for n in pack["mesh_names"]]})
        return out

    def export_poses(self, data_root_path, object_code_list, hand_name, energy_name="graspqp", suffix="None",
                     grasp_type=None):
        """writes <root>/<object>/grasp_predictions/<hand>/<n>_contacts/<energy>/<default|type>/<object><suffix>.dexgrasp.pt
        (fit.py:203-221, 288-299); -> list of paths"""
        torch = self.torch
        paths = []
        for code, data in zip(object_code_list, self.grasp_data(grasp_type)):
            d = os.path.join(data_root_path, code, "grasp_predictions", hand_name, f"{self.eng.n_c}_contacts", energy_name,
                             "default" if grasp_type in (None, "all") else grasp_type)
            os.makedirs(d, exist_ok=True)
            path = os.path.join(d, code + f"{suffix}.dexgrasp.pt")
            torch.save(data, path)
            paths.append(path)
        self.close()
        return paths
