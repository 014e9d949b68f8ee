"""TEST INFRASTRUCTURE — NOT PRODUCT CODE.

Restatement of the reference's two OTHER force-closure energies, dtype-generic torch:

* dexgrasp_energy : graspqp/src/graspqp/metrics/ops/dexgrasp.py:5-69 (`calc_e_fc`, `DexgraspSpanMetric.forward`)
                    E = | sum_i n_i |^2 + tw^2 | sum_i n_i x (p_i - cog) |^2 ; fit.py calls it with torque_weight 0.0
                    (the metric's default: energy.py:34-41 passes none).
* TDGOracle       : graspqp/src/graspqp/metrics/ops/tdg.py:57-239 (`GraspEnergyBase`, `TDGEnergy`, `TDGSpanMetric`):
                    contact frames (:84-107), grasp matrix with positions normalised by the 0.2 m box length (:109-127),
                    cone projection of the P = 1000 target directions (:165-206), density weights (:129-132),
                    E = 100 * mean_u (1 - cos(w(u), u)) (:208-214, :236-238).

Both are pinned against the reference's own modules run in the build container (tests/golden/metric_alt_ref.npz,
tests/golden/make_golden.py:gen_metric_alt).  The object normals enter as constants (TorchSDF returns them without a
graph, object_model.py:220), so gradients flow through the contact points only.
"""
import math

import torch


def dexgrasp_energy(contact_pts, contact_normals, cog, torque_weight=0.0):
    p = contact_pts - cog.unsqueeze(1)
    f = contact_normals.sum(1)                                    # contact_normals @ [I; ...]
    t = torque_weight * torch.cross(contact_normals, p, dim=-1).sum(1)  # n' skew(p) = n x p
    nrm = torch.sqrt((f * f).sum(-1) + (t * t).sum(-1))           # torch.norm(..., dim=[1, 2])
    return nrm * nrm


def _normalize(v):
    return v / torch.clamp(v.norm(dim=-1, p=2, keepdim=True), min=1e-12)


class TDGOracle:
    def __init__(self, directions, miu=0.2, obb_length=0.2, enable_density=True, scale=100.0):
        """directions (P,3): the unit target directions the reference draws with np.random.randn at construction
        (tdg.py:44-48, 148-150); injected here."""
        self.dirs = torch.as_tensor(directions)
        self.miu, self.obb, self.density, self.scale = miu, obb_length, enable_density, scale

    def frames(self, n):
        """tdg.py:84-107 -> (axis_0, axis_1, axis_2); the base vectors are float constants"""
        dt = n.dtype
        b1, b2 = torch.tensor([0, 1, 0], dtype=dt), torch.tensor([0, 0, 1], dtype=dt)
        proj = (n * b1).sum(-1, keepdim=True).abs()
        a1 = torch.where(proj > 0.99, b2.expand_as(n), b1.expand_as(n))
        a1 = _normalize(a1 - (a1 * n).sum(-1, keepdim=True) * n).detach()
        a1 = _normalize(a1 - (a1 * n).sum(-1, keepdim=True) * n)
        return n, a1, torch.cross(n, a1, dim=-1)

    def energy(self, contact_pts, contact_normals, cog):
        dt = contact_pts.dtype
        d = self.dirs.to(dt)                                                    # (P,3)
        rel = (contact_pts - cog.unsqueeze(1)) / self.obb                       # (B,n,3)   :118-121
        a0, a1, a2 = self.frames(contact_normals)
        frame = torch.stack([a0, a1, a2], -1)                                   # (B,n,3,3) columns = axes
        # G^T u for u = [d; 0]: the torque rows meet zeros, so it is frame^T d                              :173-175
        dF = _normalize(torch.einsum("bnki,pk->bpni", frame, d))                # (B,P,n,3)
        proj = dF[..., :1]                                                      # . (1,0,0)
        perp = dF.clone()
        perp[..., 0] = 0
        angles = torch.acos(torch.clamp(proj, min=-1, max=1))
        ba = math.atan(self.miu)
        r1, r2, r3 = angles <= ba, (angles > ba) & (angles <= math.pi / 2), angles > math.pi / 2
        pn = perp.norm(dim=-1, keepdim=True)
        e0 = torch.tensor([1, 0, 0], dtype=dt)
        help3 = pn / (pn - 2 * self.miu * torch.clamp(proj, max=0))
        help2 = e0 + self.miu * _normalize(perp)
        c1 = e0 + perp / torch.clamp(proj, min=math.cos(ba) / 2)
        c = torch.where(r1, c1, torch.where(r2, help2, help3 * help2))          # the reference multiplies by the masks
        f = torch.einsum("bnki,bpni->bpnk", frame, c)                           # world-frame force of contact n
        w = torch.cat([f, torch.cross(rel.unsqueeze(1).expand_as(f), f, dim=-1)], -1)  # (B,P,n,6)   :199
        if self.density:
            cos = (contact_normals.unsqueeze(-2) * contact_normals.unsqueeze(-3)).sum(-1)
            rho = (1 / torch.clamp(torch.clamp(cos, min=0).sum(-1), min=1e-4)).detach()      # :129-132
            W = (w * rho.unsqueeze(1).unsqueeze(-1)).sum(2)
        else:
            W = w.sum(2)
        u6 = torch.cat([d, torch.zeros_like(d)], -1).unsqueeze(0)
        cos_wt = (_normalize(W) * u6).sum(-1)
        return self.scale * (1 - cos_wt).mean(-1)
