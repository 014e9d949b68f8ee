"""TEST INFRASTRUCTURE — NOT PRODUCT CODE.

Restatement of the per-iteration loop body of scripts/fit.py:399-458 with INJECTED noise:
* ObjectOracle.cal_distance: graspqp/src/graspqp/core/object_model.py:186-255 (TORCHSDF branch), cog :64-68
* calculate_energy: graspqp/src/graspqp/core/energy.py:6-66 (method "gendexgrasp": fit.py passes use_gendexgrasp)
* MalaStarOracle: graspqp/src/graspqp/core/optimizer.py:152-347
* loop glue (z-score, weighted sum, masked commit): scripts/fit.py:381-396, 403-406, 434-458

The reference draws `torch.rand(B,n_c)`, `torch.randint(C, [switch.sum()])` and `torch.rand(B)`; here they are the
arguments u_switch (B,n_c), idx_new (B,n_c) (only entries under the switch mask are used) and u_accept (B).
"""
import numpy as np
import torch

from .hand import HandOracle, compute_sdf
from .metrics_alt import TDGOracle, dexgrasp_energy
from .qp import graspqp_energy

WEIGHTS = dict(E_dis=100.0, E_fc=1.0, E_pen=100.0, E_spen=10.0, E_joints=1.0)  # fit.py argparse defaults :90-112


class ObjectOracle:
    def __init__(self, face_verts_list, surface_points_list, batch_size_each, dtype=torch.float32, device="cpu",
                 sdf=None):
        self.face_verts_list = face_verts_list  # list of (F,3,3) float32 numpy
        self.sdf = sdf if sdf is not None else compute_sdf
        if sdf is not None:  # GPU baseline leg: meshes as device tensors for the reference's torchsdf.compute_sdf
            self.face_verts_list = [torch.as_tensor(np.asarray(f, np.float32).reshape(-1, 3, 3)).to(device).contiguous()
                                    for f in face_verts_list]
        self.batch_size_each = batch_size_each
        self.dtype = dtype
        sp = torch.stack([torch.as_tensor(s, dtype=dtype) for s in surface_points_list], 0).to(device)
        self.surface_points_tensor = sp.repeat_interleave(batch_size_each, dim=0)
        self.object_scale = 1.0  # object_model.py:60 — always 1.0
        self.cog = self.surface_points_tensor.mean(1)

    def cal_distance(self, x):
        n_points = x.shape[1]
        xo = x.reshape(-1, self.batch_size_each * n_points, 3) / self.object_scale
        distance, normals = [], []
        for i, fv in enumerate(self.face_verts_list):
            dis, sg, normal, _ = self.sdf(xo[i], fv)
            dis = torch.sqrt(dis + 1e-8) * (-sg)
            distance.append(dis)
            normals.append(normal * sg.unsqueeze(1))
        distance = (torch.stack(distance) * self.object_scale).reshape(-1, n_points)
        return distance, torch.stack(normals).reshape(-1, n_points, 3)


def calculate_energy(hand, obj, svd_gain=0.1, friction=0.2, max_limit=20.0, energy_type="graspqp", tdg=None,
                     dexgrasp_torque_weight=0.0, energy_names=()):
    """energy_type: graspqp | dexgrasp | tdg (scripts/fit.py:344-356); tdg = a metrics_alt.TDGOracle"""
    distance, contact_normal = obj.cal_distance(hand.contact_points)
    E_dis = ((1 - torch.sum((-contact_normal) * hand.contact_normals, dim=-1)).exp() * distance.abs()).sum(-1)
    info = {}
    if energy_type == "graspqp":
        E_fc, lam, info = graspqp_energy(hand.contact_points, contact_normal, obj.cog, friction=friction,
                                         max_limit=max_limit, svd_gain=svd_gain)
    elif energy_type == "dexgrasp":
        E_fc = dexgrasp_energy(hand.contact_points, contact_normal, obj.cog, dexgrasp_torque_weight)
    elif energy_type == "tdg":
        E_fc = tdg.energy(hand.contact_points, contact_normal, obj.cog)
    else:
        raise ValueError(energy_type)
    th = hand.hand_pose[:, 9:]
    E_joints = torch.sum((th > hand.joints_upper) * (th - hand.joints_upper), dim=-1) + torch.sum(
        (th < hand.joints_lower) * (hand.joints_lower - th), dim=-1)
    d = hand.cal_distance(obj.surface_points_tensor * obj.object_scale)
    d = d.clone()
    d[d <= 0] = 0
    E_pen = d.sum(-1)
    E_spen = hand.self_penetration()
    losses = dict(E_dis=E_dis, E_fc=E_fc, E_joints=E_joints, E_pen=E_pen, E_spen=E_spen)
    if "E_prior" in energy_names:  # energy.py:68-74
        forward_axis = (hand.global_rotation @ hand.grasp_axis.view(1, -1, 1)).view(-1, 3)
        axis_prior = torch.tensor([0, 0, -1], dtype=hand.dtype).view(1, 3)
        losses["E_prior"] = 1 - torch.sum(forward_axis * axis_prior, dim=-1)
    if "E_wall" in energy_names:  # energy.py:75-77
        losses["E_wall"] = hand.get_surface_points()[..., -1].clamp(max=0.0).abs().sum(-1)
    return losses, info


def total_energy(losses, weights=WEIGHTS):
    e = 0
    for k, v in losses.items():
        e = e + weights[k] * v
    return e


class MalaStarOracle:
    def __init__(self, hand, switch_possibility=0.4, starting_temperature=18.0, temperature_decay=0.95,
                 annealing_period=30, step_size=0.005, stepsize_period=50, mu=0.98, clip_grad=False):
        self.hand = hand
        self.clip_grad = clip_grad
        self.switch_possibility = switch_possibility
        dt = hand.dtype
        self.starting_temperature = torch.tensor(starting_temperature, dtype=dt)
        self.temperature_decay = torch.tensor(temperature_decay, dtype=dt)
        self.annealing_period, self.step_size_period = annealing_period, stepsize_period
        self.step_size = torch.tensor(step_size, dtype=dt)
        self.mu = torch.tensor(mu, dtype=dt)
        B = hand.hand_pose.shape[0]
        self.step = torch.zeros(B, dtype=torch.long)
        self.ema = torch.zeros(B, hand.n_dofs + 9, dtype=dt)

    def try_step(self, u_switch, idx_new, grad2_mean=None):
        h = self.hand
        s = self.step_size * self.temperature_decay ** torch.div(self.step, self.step_size_period, rounding_mode="floor")
        step_size = torch.zeros_like(h.hand_pose) + s[..., None]
        if self.clip_grad:  # optimizer.py:211-213
            gradient = h.hand_pose.grad.clip(min=-100, max=100)
            gradient[torch.isnan(gradient)] = 0
        else:
            gradient = h.hand_pose.grad
        grad = (gradient ** 2).mean(0)
        if grad2_mean is not None:  # sub-batch checks: the statistic of the FULL batch this subset was taken from
            grad = torch.as_tensor(grad2_mean, dtype=h.dtype)
        self.ema = self.mu * grad + (1 - self.mu) * self.ema
        self.ema[torch.isnan(self.ema)] = 0
        hand_pose = h.hand_pose - step_size * gradient / (torch.sqrt(self.ema) + 1e-6)
        hand_pose = hand_pose.detach()
        if hand_pose.isnan().any():  # optimizer.py:242-244
            hand_pose[hand_pose.isnan().any(dim=-1)] = 0
        switch_mask = u_switch < self.switch_possibility
        idx = h.contact_point_indices.clone()
        idx[switch_mask] = idx_new[switch_mask]
        self.old = dict(pose=h.hand_pose, idx=h.contact_point_indices, grad=h.hand_pose.grad)
        h.set_parameters(hand_pose.requires_grad_(True), idx)
        self.step += 1
        return s

    def reset_envs(self, mask):  # optimizer.py:275-287
        h = self.hand
        self.step[mask] = 0
        self.ema[mask] = 0
        with torch.no_grad():
            self.old["pose"][mask] = h.hand_pose[mask]
        self.old["idx"][mask] = h.contact_point_indices[mask]
        self.old["grad"][mask] = 0 * self.old["grad"][mask]

    def accept_step(self, energy, new_energy, u_accept, z_score, reset_mask=None, after_reset=False):
        """after_reset: the PREVIOUS iteration was a reset iteration — see AFTER_RESET below"""
        h = self.hand
        temperature = self.starting_temperature * self.temperature_decay ** torch.div(
            self.step, self.annealing_period, rounding_mode="floor")
        proba = torch.distributions.normal.Normal(0.0, 1.0).cdf(z_score.detach())
        temperature = temperature * (1 + proba)
        accept = u_accept < torch.exp((energy - new_energy) / temperature)
        if reset_mask is not None:
            accept[reset_mask] = True  # optimizer.py:315-316
        with torch.no_grad():
            reject = ~accept
            h.hand_pose[reject] = self.old["pose"][reject]
            h.contact_point_indices[reject] = self.old["idx"][reject]
            old_grad = self.old["grad"] + h.hand_pose.grad if after_reset else self.old["grad"]
            h.hand_pose.grad[reject] = old_grad[reject]
        return accept, temperature


AFTER_RESET = """A behaviour of the reference that both optimisers share, restated explicitly (the oracle detaches its proposals):
try_step builds the proposal from hand_model.hand_pose WITHOUT detaching it (optimizer.py:88-90, 240) and
HandModel.set_parameters keeps `hand_pose.clone()` (hand_model.py:855), so the proposal's graph runs back into the previous
pose with an identity Jacobian and the backward pass of the new energy also delivers the new gradient to the previous pose.
Normally that pose is a non-leaf whose retained `.grad` is REPLACED by a new tensor (old + new), which `old_grad_hand_pose`
(an alias of the old tensor object, :104 / :268) does not see.  But a reset iteration builds the pose with torch.where under
no_grad and requires_grad = True (hand_model.py:845-850): a LEAF, into whose `.grad` autograd accumulates IN PLACE.  In the
iteration right after a reset iteration the alias therefore holds old + new gradient, and that is what rejected rows get
back (:136 / :338).  Seen in the reference runs of tests/golden/fit_alt_ref.npz (make_golden.gen_fit_alt)."""


class AnnealingOracle:
    """AnnealingDexGraspNet, graspqp/src/graspqp/core/optimizer.py:11-149 (--optimizer dexgraspnet): one RMS accumulator
    and one step counter for the batch, no clipping / NaN handling, temperature without the z-score factor, resets do not
    touch the optimiser (reset_envs :148-149 is `pass`) and are not force-accepted.

    See AFTER_RESET for what `after_reset` restates."""

    def __init__(self, hand, switch_possibility=0.5, starting_temperature=18.0, temperature_decay=0.95,
                 annealing_period=30, step_size=0.005, stepsize_period=50, mu=0.98, clip_grad=False):
        self.hand = hand
        self.switch_possibility = switch_possibility
        dt = hand.dtype
        self.starting_temperature = torch.tensor(starting_temperature, dtype=dt)
        self.temperature_decay = torch.tensor(temperature_decay, dtype=dt)
        self.annealing_period, self.step_size_period = annealing_period, stepsize_period
        self.step_size = torch.tensor(step_size, dtype=dt)
        self.mu = torch.tensor(mu, dtype=dt)
        self.step = 0
        self.ema = torch.zeros(hand.n_dofs + 9, dtype=dt)

    def try_step(self, u_switch, idx_new, grad2_mean=None):
        h = self.hand
        s = self.step_size * self.temperature_decay ** (self.step // self.step_size_period)
        g2 = (h.hand_pose.grad ** 2).mean(0) if grad2_mean is None else torch.as_tensor(grad2_mean, dtype=h.dtype)
        self.ema = self.mu * g2 + (1 - self.mu) * self.ema
        hand_pose = (h.hand_pose - s * h.hand_pose.grad / (torch.sqrt(self.ema) + 1e-6)).detach()
        switch_mask = u_switch < self.switch_possibility
        idx = h.contact_point_indices.clone()
        idx[switch_mask] = idx_new[switch_mask]
        self.old = dict(pose=h.hand_pose, idx=h.contact_point_indices, grad=h.hand_pose.grad)
        h.set_parameters(hand_pose.requires_grad_(True), idx)
        self.step += 1
        return s

    def reset_envs(self, mask):
        pass

    def accept_step(self, energy, new_energy, u_accept, z_score=None, reset_mask=None, after_reset=False):
        h = self.hand
        temperature = self.starting_temperature * self.temperature_decay ** (self.step // self.annealing_period)
        accept = u_accept < torch.exp((energy - new_energy) / temperature)
        with torch.no_grad():
            reject = ~accept
            h.hand_pose[reject] = self.old["pose"][reject]
            h.contact_point_indices[reject] = self.old["idx"][reject]
            old_grad = self.old["grad"] + h.hand_pose.grad if after_reset else self.old["grad"]
            h.hand_pose.grad[reject] = old_grad[reject]
        return accept, temperature


class FitLoopOracle:
    """One object of this class = the state of scripts/fit.py between iterations."""

    def __init__(self, pack, face_verts_list, surface_points_list, batch_size_each, hand_pose0, idx0,
                 dtype=torch.float32, energy_type="graspqp", optimizer="mala_star", tdg_directions=None,
                 dexgrasp_torque_weight=0.0, weights=None, hand_surface=None, **opt):
        """weights: overrides of WEIGHTS; terms E_prior / E_wall are evaluated when their weight is > 0 (fit.py:373);
        hand_surface = (points, link) for E_wall"""
        self.hand = HandOracle(pack, dtype)
        if hand_surface is not None:
            self.hand.set_surface_points(*hand_surface)
        self.weights = dict(WEIGHTS)
        self.weights.update(weights or {})
        self.obj = ObjectOracle(face_verts_list, surface_points_list, batch_size_each, dtype)
        self.B = batch_size_each
        self.hand.set_parameters(torch.as_tensor(hand_pose0, dtype=dtype).clone().requires_grad_(True),
                                 torch.as_tensor(idx0).long())
        self.opt = (MalaStarOracle if optimizer == "mala_star" else AnnealingOracle)(self.hand, **opt)
        self.ekw = dict(energy_type=energy_type, dexgrasp_torque_weight=dexgrasp_torque_weight,
                        tdg=TDGOracle(tdg_directions) if energy_type == "tdg" else None,
                        energy_names=[k for k in ("E_prior", "E_wall") if self.weights.get(k, 0.0) > 0.0])
        self.losses, _ = calculate_energy(self.hand, self.obj, **self.ekw)
        self.energy = total_energy(self.losses, self.weights)
        self.energy.sum().backward()
        self.first_grad = self.hand.hand_pose.grad.clone()
        self.hand.hand_pose.grad.zero_()  # fit.py:396
        self.energy = self.energy.detach()
        self.losses = {k: v.detach().clone() for k, v in self.losses.items()}

    def load_state(self, pose, idx, grad, energy, ema, step):
        """Teacher forcing: take over the state between two iterations from another run of the same loop."""
        dt = self.hand.dtype
        self.hand.set_parameters(torch.as_tensor(pose, dtype=dt).clone().requires_grad_(True), torch.as_tensor(idx).long())
        self.hand.hand_pose.grad = torch.as_tensor(grad, dtype=dt).clone()
        self.energy = torch.as_tensor(energy, dtype=dt).clone()
        self.opt.ema = torch.as_tensor(ema, dtype=dt).clone()
        self.opt.step = int(step) if isinstance(self.opt, AnnealingOracle) else torch.as_tensor(step).long().clone()

    def iterate(self, u_switch, idx_new, u_accept, reset=None, z_score_threshold=1.0, grad2_mean=None, z=None,
                qp_stop=None):
        """reset = (pose_init (Bt,D), idx_init (Bt,n_c), mask or None): a reset iteration, scripts/fit.py:408-422 with
        initialize_convex_hull's result injected (initializations.py:190-193 -> set_parameters(env_mask)).
        grad2_mean / z / qp_stop: batch-wide statistics of the full batch when this oracle holds a SUBSET of it."""
        from . import qp as _qp
        s = self.opt.try_step(u_switch, idx_new, grad2_mean)
        if z is None:
            Eb = self.energy.view(-1, self.B)
            z = ((Eb - Eb.mean(-1).unsqueeze(-1)) / Eb.std(-1).unsqueeze(-1)).view(-1)
        else:
            z = torch.as_tensor(z, dtype=self.hand.dtype)
        reset_mask = None
        if reset is not None:
            pose_init, idx_init, mask = reset
            reset_mask = (z > z_score_threshold) if mask is None else torch.as_tensor(mask).bool()
            hp = torch.as_tensor(pose_init, dtype=self.hand.dtype).clone().requires_grad_(True)
            self.hand.set_parameters(hp, torch.as_tensor(idx_init).long(), env_mask=reset_mask)
            self.opt.reset_envs(reset_mask)
        _qp.FORCE_STOP = qp_stop
        try:
            new_losses, info = calculate_energy(self.hand, self.obj, **self.ekw)
        finally:
            _qp.FORCE_STOP = None
        new_energy = total_energy(new_losses, self.weights)
        new_energy.sum().backward()
        out = dict(energy_before=self.energy.detach().clone(), new_energy=new_energy.detach().clone(), new_grad=self.hand.hand_pose.grad.clone(),
                   new_losses={k: v.detach().clone() for k, v in new_losses.items()}, z=z, s=s,
                   proposal=self.hand.hand_pose.detach().clone(), idx_prop=self.hand.contact_point_indices.clone())
        with torch.no_grad():
            accept, t = self.opt.accept_step(self.energy, new_energy.detach(), u_accept, z, reset_mask,
                                             after_reset=getattr(self, "prev_reset", False))
            self.prev_reset = reset is not None
            self.energy[accept] = new_energy.detach()[accept]
            for k, v in new_losses.items():
                self.losses[k][accept] = v.detach()[accept]
        out.update(accept=accept, temperature=t, reset_mask=reset_mask)
        return out
