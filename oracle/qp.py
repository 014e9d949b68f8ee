"""TEST INFRASTRUCTURE — NOT PRODUCT CODE.

Restatement of the force-closure metric for energy_type=graspqp:
* friction cone + grasp matrix + metric: graspqp/src/graspqp/metrics/ops/span.py:263-295, 313-415
* wrapper / E_fc: graspqp/src/graspqp/metrics/ops/registry.py:31-89
* QP construction: graspqp/src/graspqp/metrics/solver/qp_solver.py:60-134
* the QP solve itself happens in `qpth.qp.QPFunction(verbose=False, maxIter=12, eps=5e-2)` (qp_solver.py:8,125).
  qpth==0.0.18 (graspqp/pyproject.toml:40) is NOT vendored under /root/reference and is not installed here.  What
  follows restates its published algorithm (qpth/qp.py QPFunction + qpth/solvers/pdipm/batch.py: pre_factor_kkt,
  factor_kkt, solve_kkt (LU_PARTIAL path), forward, get_step) from the package's public source as recalled:
  batched primal-dual interior point (Mehrotra predictor-corrector), KKT reduced by Schur complement
  S = G Q^-1 G^T + diag(s/z); start = solve with d=1 then shift s,z to >= 1; residual
  |r_x| + |r_z| + nineq*mu; each element keeps its best-residual iterate; the loop exits when the BATCH MAX of the
  best residuals < eps, when no element improved for 3 consecutive iterations, or after maxIter; backward solves one
  more KKT system with d = clamp(lam)/clamp(slack) and returns dQ = (dx x^T + x dx^T)/2.
  PARITY UNPINNED for the truncated iterate and the backward; the only reference-held known answer
  (graspqp/tests/metrics/test_solver.py:5-26, converged residual ~ 0, atol 1e-4) is checked in tests/test_oracle_qp.py,
  together with the converged optimum from scipy.optimize.lsq_linear (the reference's own ScipyLsqSolver,
  metrics/solver/scipy_solver.py:64-125).
"""
import math

import torch


def _get_step(v, dv):
    a = -v / dv
    a = a.clone()
    a[dv >= 0] = max(1.0, a.max().item())
    return a.min(1)[0]


class _KKT:
    """LU_PARTIAL path specialised to neq = 0 (the only case on this path)."""

    def __init__(self, Q, G):
        self.Q, self.G = Q, G
        self.Q_LU = torch.linalg.lu_factor(Q)
        # R = G Q^-1 G^T
        G_invQ_GT = torch.bmm(G, torch.linalg.lu_solve(*self.Q_LU, G.transpose(1, 2)))
        self.R = G_invQ_GT

    def factor(self, d):
        T = self.R.clone()
        idx = torch.arange(T.shape[1], device=T.device)
        T[:, idx, idx] += 1.0 / d
        self.S_LU = torch.linalg.lu_factor(T)

    def solve(self, d, rx, rs, rz):
        invQ_rx = torch.linalg.lu_solve(*self.Q_LU, rx.unsqueeze(2)).squeeze(2)
        h = torch.bmm(invQ_rx.unsqueeze(1), self.G.transpose(1, 2)).squeeze(1) + rs / d - rz
        w = -torch.linalg.lu_solve(*self.S_LU, h.unsqueeze(2)).squeeze(2)
        g1 = -rx - torch.bmm(w.unsqueeze(1), self.G).squeeze(1)
        g2 = -rs - w
        dx = torch.linalg.lu_solve(*self.Q_LU, g1.unsqueeze(2)).squeeze(2)
        return dx, g2 / d, w


# Sub-batch checks (tests/cases.py:check_benched_state): qpth's exit test is batch-wide, so a SUBSET of a large batch
# can only be reproduced if the loop leaves at the iteration the FULL batch left at.  FORCE_STOP = that iteration
# (max_iter = never left early); None = qpth's own test on the batch at hand.
FORCE_STOP = None


def pdipm_forward(Q, p, G, h, eps=5e-2, maxIter=12, notImprovedLim=3, trace=None):
    """Returns x, lam (z), slack (s), kkt.  `trace`, if a list, receives per-iteration residual vectors."""
    nB, nineq, nz = G.shape[0], G.shape[1], G.shape[2]
    kkt = _KKT(Q, G)
    d = torch.ones(nB, nineq, dtype=Q.dtype, device=Q.device)
    kkt.factor(d)
    x, s, z = kkt.solve(d, p, torch.zeros(nB, nineq, dtype=Q.dtype, device=Q.device), -h)
    for v in (s, z):
        M = torch.min(v, 1)[0].view(-1, 1).repeat(1, nineq)
        I = M < 0
        v[I] -= M[I] - 1
    best = None
    nNotImproved = 0
    for i in range(maxIter):
        rx = torch.bmm(z.unsqueeze(1), G).squeeze(1) + torch.bmm(x.unsqueeze(1), Q.transpose(1, 2)).squeeze(1) + p
        rs = z
        rz = torch.bmm(x.unsqueeze(1), G.transpose(1, 2)).squeeze(1) + s - h
        mu = torch.abs((s * z).sum(1) / nineq)
        resids = torch.norm(rz, 2, 1) + torch.norm(rx, 2, 1) + nineq * mu
        if trace is not None:
            trace.append(resids.clone())
        d = z / s
        kkt.factor(d)
        if best is None:
            best = dict(resids=resids.clone(), x=x.clone(), z=z.clone(), s=s.clone())
            nNotImproved = 0
        else:
            I = resids < best["resids"]
            nNotImproved = 0 if I.sum() > 0 else nNotImproved + 1
            best["resids"][I] = resids[I]
            best["x"][I] = x[I]
            best["z"][I] = z[I]
            best["s"][I] = s[I]
        if FORCE_STOP is not None:
            if i == FORCE_STOP:
                return best["x"], best["z"], best["s"], kkt
        elif nNotImproved == notImprovedLim or best["resids"].max() < eps or mu.min() > 1e32:
            return best["x"], best["z"], best["s"], kkt
        dx_aff, ds_aff, dz_aff = kkt.solve(d, rx, rs, rz)
        alpha = torch.min(torch.min(_get_step(z, dz_aff), _get_step(s, ds_aff)), torch.ones(nB, dtype=Q.dtype, device=Q.device))
        a = alpha.view(-1, 1)
        t3 = torch.sum((s + a * ds_aff) * (z + a * dz_aff), 1)
        t4 = torch.sum(s * z, 1)
        sig = (t3 / t4) ** 3
        rs2 = ((-mu * sig).view(-1, 1) + ds_aff * dz_aff) / s
        zx, zi = torch.zeros(nB, nz, dtype=Q.dtype, device=Q.device), torch.zeros(nB, nineq, dtype=Q.dtype, device=Q.device)
        dx_cor, ds_cor, dz_cor = kkt.solve(d, zx, rs2, zi)
        dx, ds, dz = dx_aff + dx_cor, ds_aff + ds_cor, dz_aff + dz_cor
        alpha = torch.min(0.999 * torch.min(_get_step(z, dz), _get_step(s, ds)), torch.ones(nB, dtype=Q.dtype, device=Q.device))
        a = alpha.view(-1, 1)
        x = x + a * dx
        s = s + a * ds
        z = z + a * dz
    return best["x"], best["z"], best["s"], kkt


class QPFunction(torch.autograd.Function):
    """qpth.qp.QPFunction for neq = 0, shared G (nineq,nz), per-element h; grads for Q and p only (G, h constant)."""

    @staticmethod
    def forward(ctx, Q, p, G, h):
        nB = Q.shape[0]
        Gb = G.unsqueeze(0).expand(nB, -1, -1).contiguous()
        x, lam, slack, kkt = pdipm_forward(Q.detach(), p.detach(), Gb, h)
        ctx.kkt, ctx.lam, ctx.slack = kkt, lam, slack
        ctx.save_for_backward(x)
        return x

    @staticmethod
    def backward(ctx, dl_dx):
        (x,) = ctx.saved_tensors
        d = torch.clamp(ctx.lam, min=1e-8) / torch.clamp(ctx.slack, min=1e-8)
        ctx.kkt.factor(d)
        z0 = torch.zeros_like(ctx.lam)
        dx, _, _ = ctx.kkt.solve(d, dl_dx, z0, z0)
        dQ = 0.5 * (dx.unsqueeze(2) * x.unsqueeze(1) + x.unsqueeze(2) * dx.unsqueeze(1))
        return dQ, dx, None, None


def lsq_solve(A, b, min_bound, max_bound):
    """qp_solver.py:60-134 (SQPLsqSolver.solve) for A (B,6,n), b (B,6). Returns value (B), x (B,n)."""
    n = A.shape[-1]
    kw = dict(dtype=A.dtype, device=A.device)
    u = torch.ones(A.shape[0], n, **kw) * max_bound
    l = torch.ones(A.shape[0], n, **kw) * min_bound
    Q = A.mT @ A
    Q = Q + torch.eye(n, **kw).unsqueeze(0) * 1e-4
    p = (-A.mT @ b[..., None]).squeeze(-1)
    G = torch.cat([torch.eye(n, **kw), -torch.eye(n, **kw)], dim=-2)
    h = torch.cat([u, -l], dim=-1)
    x = QPFunction.apply(Q, p, G, h)
    value = 0.5 * torch.sum((b - (A @ x.unsqueeze(-1)).squeeze(-1)).pow(2), -1)
    return value, x


def friction_cone(normals, mu=0.2):
    """span.py:263-295, n_cone_vecs == 4."""
    b1 = torch.ones_like(normals) / math.sqrt(3)
    dot_product = torch.sum(b1 * normals, dim=-1) * 1 / (normals.norm(dim=-1) + 1e-6)
    b1 = b1.clone()
    b1[..., 1] -= 2 * (dot_product > 0.9).to(normals.dtype)
    t1 = torch.linalg.cross(normals, b1)
    t2 = torch.linalg.cross(normals, t1)
    c = math.sqrt(1 - mu ** 2)
    tb = [mu * t1 + c * normals, mu * t2 + c * normals, -mu * t1 + c * normals, -mu * t2 + c * normals]
    return torch.stack(tb, dim=-2).flatten(-3, -2) / len(tb)


def grasp_matrix(contact_pts, contact_normals, cog, mu=0.2, torque_weight=5.0):
    """span.py:341-346 -> F (B,6,4*n_c)."""
    r = (contact_pts - cog.unsqueeze(1)).repeat_interleave(4, dim=-2)
    f = friction_cone(contact_normals, mu)
    tau = torch.linalg.cross(r, f) * torque_weight
    return torch.cat([f, tau], dim=-1).mT


def graspqp_energy(contact_pts, contact_normals, cog, friction=0.2, max_limit=20.0, svd_gain=0.1, torque_weight=5.0,
                   values_gain=2.0):
    """registry.py:31-89 with span.py:313-415: returns E_fc (B), per-contact force sums (B,n_c), and pieces."""
    F = grasp_matrix(contact_pts, contact_normals, cog, friction, torque_weight)
    b = torch.zeros(F.shape[0], 6, dtype=F.dtype, device=F.device)
    value, x = lsq_solve(F, b, 1.0, max_limit + 1.0)
    svd = torch.linalg.svdvals(F).prod(-1) ** (1.0 / 6.0)
    E = values_gain * (value + 1e-2) * torch.exp(-svd_gain * svd)
    return E, x.view(x.shape[0], -1, 4).sum(-1), dict(F=F, value=value, svd=svd, x=x)
