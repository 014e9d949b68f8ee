"""TEST INFRASTRUCTURE — NOT PRODUCT CODE.

Restatement of HandModel.joint_entropy / pose_entropy (graspqp/src/graspqp/core/hand_model.py:698-760), the statistics
scripts/fit.py:462-471 logs with --log_entropy.  PARITY UNPINNED: the methods live on HandModel, whose module imports trimesh
and pytorch3d (absent), and pose_entropy calls roma.rotmat_to_rotvec (absent: restated from roma 1.5's rotmat_to_unitquat /
unitquat_to_rotvec — shortest arc, w >= 0).  torch.histc itself is the real one.
"""
import numpy as np
import torch

from .hand import special_gramschmidt


def _hist_entropy(x, lo, hi, n_bins=32):
    counts = x.histc(n_bins, float(lo), float(hi))
    dist = counts / counts.sum()
    logs = torch.log(torch.where(dist > 0, dist, torch.ones_like(dist)))
    return -(dist * logs).sum()


def joint_entropy(hand_pose, joints_lower, joints_upper):
    th = hand_pose[:, 9:]
    e = 0
    for j in range(th.shape[1]):
        e = e + _hist_entropy(th[:, j], joints_lower[j], joints_upper[j]) / th.shape[1]
    return e


def rotmat_to_rotvec(R):
    """roma.rotmat_to_rotvec: through the unit quaternion (xyzw) with w >= 0; angle = 2 atan2(|v|, w)"""
    from scipy.spatial.transform import Rotation
    q = torch.as_tensor(Rotation.from_matrix(R.detach().double().numpy()).as_quat(), dtype=R.dtype)  # xyzw
    q = torch.where(q[:, 3:] < 0, -q, q)
    vn = q[:, :3].norm(dim=1)
    angle = 2 * torch.atan2(vn, q[:, 3])
    small = angle.abs() <= 1e-3
    scale = torch.where(small, 2 + angle ** 2 / 12 + 7 * angle ** 4 / 2880, angle / torch.sin(angle / 2))
    return scale[:, None] * q[:, :3]


def pose_entropy(hand_pose):
    t = sum(_hist_entropy(hand_pose[:, i], -0.1, 0.1) / 3 for i in range(3))
    R = special_gramschmidt(torch.stack([hand_pose[:, 3:6], hand_pose[:, 6:9]], -1))
    rv = rotmat_to_rotvec(R)
    r = torch.norm(rv, dim=-1)
    theta = torch.acos(rv[:, 2] / r)
    phi = torch.sign(rv[:, 1]) * torch.acos(rv[:, 0] / torch.norm(rv[:, :2], dim=-1))
    sph = torch.stack([r, theta, phi], -1)
    limits = [(0, np.pi), (0, np.pi), (-np.pi, np.pi)]
    rot = sum(_hist_entropy(sph[:, i], *limits[i]) / 3 for i in range(3))
    return t, rot
