"""MalaStar — graspqp/src/graspqp/core/optimizer.py:152-347 — and AnnealingDexGraspNet — :11-149 — with the reference's
constructors and methods (try_step / accept_step / zero_grad / reset_envs), over the csrc/gqp_mala.cu kernels.  It reads and writes the same
hand_model attributes the reference does (hand_pose(.grad), contact_point_indices, ...).  Noise comes from Philox4x32-10
on the device (seed, iteration, row), or from `inject_noise(...)` for parity runs."""
import ctypes

import torch

from . import _lib


class MalaStar:
    def __init__(self, hand_model, switch_possibility=0.5, starting_temperature=18, temperature_decay=0.95,
                 annealing_period=30, step_size=0.005, stepsize_period=50, mu=0.98, device="cuda", global_ema=False,
                 clip_grad=False, batch_size=-1, seed=1):
        self.hand_model, self.device, self.batch_size = hand_model, torch.device(device), batch_size
        self.params = _lib.GqpMalaParams(switch_possibility=switch_possibility, starting_temperature=starting_temperature,
                                         temperature_decay=temperature_decay, step_size=step_size, mu=mu,
                                         annealing_period=annealing_period, stepsize_period=stepsize_period,
                                         clip_grad=int(clip_grad))
        B, D = hand_model.hand_pose.shape
        dev = hand_model.hand_pose.device
        self.step = torch.zeros(B, dtype=torch.int32, device=dev)
        self.ema_grad_hand_pose = torch.zeros(B, D, device=dev)
        self.seed, self.iteration = int(seed), 0
        self._noise = None
        self.old_hand_pose = self.old_contact_point_indices = self.old_grad_hand_pose = None
        self._g2, self._z = torch.zeros(D, device=dev), torch.zeros(B, device=dev)
        self._scratch = torch.zeros((B + 63) // 64 * D, device=dev)
        self._one = torch.ones(1, device=dev)
        self._prev_reset = self._reset_now = False

    def inject_noise(self, u_switch, idx_rand, u_accept):
        """parity runs: the arrays torch.rand / torch.randint would have produced (optimizer.py:252-257, :305)"""
        self._noise = (u_switch.float().contiguous(), idx_rand.to(torch.int32).contiguous(), u_accept.float().contiguous())

    def _st(self):
        return ctypes.c_void_p(torch.cuda.current_stream(self._g2.device).cuda_stream)

    def try_step(self):
        lib, hm = _lib.lib(), self.hand_model
        pose = hm.hand_pose.detach().contiguous()
        grad = hm.hand_pose.grad if hm.hand_pose.grad is not None else torch.zeros_like(pose)
        grad = grad.contiguous()
        idx = hm.contact_point_indices.to(torch.int32).contiguous()
        B, D = pose.shape
        n_c = idx.shape[1]
        self.iteration += 1
        dummy_e = self._z  # the z-score is not needed here
        _lib.check(lib.gqp_mala_stats(grad.data_ptr(), dummy_e.data_ptr(), 1, B, D, self.params.clip_grad,
                                      self._g2.data_ptr(), torch.empty_like(self._z).data_ptr(),
                                      self._scratch.data_ptr(), self._st()))
        pose_new, idx_new = torch.empty_like(pose), torch.empty_like(idx)
        s = torch.empty(B, device=pose.device)
        us, ir = (self._noise[0], self._noise[1]) if self._noise is not None else (None, None)
        _lib.check(lib.gqp_mala_propose(ctypes.byref(self.params), pose.data_ptr(), grad.data_ptr(), self._g2.data_ptr(),
                                        self.ema_grad_hand_pose.data_ptr(), self.step.data_ptr(), idx.data_ptr(), B, D,
                                        n_c, hm.n_contact_candidates, us.data_ptr() if us is not None else None,
                                        ir.data_ptr() if ir is not None else None, self.seed, self.iteration,
                                        pose_new.data_ptr(), idx_new.data_ptr(), s.data_ptr(), self._st()))
        self.old_hand_pose, self.old_contact_point_indices = hm.hand_pose, hm.contact_point_indices  # :259-266
        self.old_grad_hand_pose = grad
        hm.set_parameters(pose_new.requires_grad_(True), idx_new.to(hm.contact_point_indices.dtype))
        return s

    def accept_step(self, energy, new_energy, reset_mask=None, z_score=None, z_score_threshold=2.0):
        lib, hm = _lib.lib(), self.hand_model
        B, D = hm.hand_pose.shape
        n_c = hm.contact_point_indices.shape[1]
        e_old = energy.detach().float().contiguous().clone()
        e_new = new_energy.detach().float().contiguous()
        z = z_score.detach().float().contiguous() if z_score is not None else None
        rm = reset_mask.to(torch.uint8).contiguous() if reset_mask is not None else None
        ua = self._noise[2] if self._noise is not None else None
        acc = torch.empty(B, dtype=torch.uint8, device=e_new.device)
        T = torch.empty(B, device=e_new.device)
        # rows are committed into copies of the OLD state: accepted rows take the proposal, rejected rows keep it
        pose = self.old_hand_pose.detach().clone().contiguous()
        idx = self.old_contact_point_indices.to(torch.int32).clone().contiguous()
        grad = self.old_grad_hand_pose.clone().contiguous()
        pose_new = hm.hand_pose.detach().contiguous()
        idx_new = hm.contact_point_indices.to(torch.int32).contiguous()
        grad_new = (hm.hand_pose.grad if hm.hand_pose.grad is not None else torch.zeros_like(pose_new)).contiguous()
        losses = torch.empty(1, B, device=e_new.device)
        p = lambda t: t.data_ptr() if t is not None else None
        _lib.check(lib.gqp_mala_accept(ctypes.byref(self.params), B, D, n_c, 1, self._one.data_ptr(), e_new.data_ptr(),
                                       p(z), p(rm), p(ua), self.seed, self.iteration, self.step.data_ptr(),
                                       pose_new.data_ptr(), idx_new.data_ptr(), grad_new.data_ptr(), pose.data_ptr(),
                                       idx.data_ptr(), grad.data_ptr(), e_old.data_ptr(), losses.data_ptr(), None,
                                       acc.data_ptr(), T.data_ptr(), int(self._prev_reset), self._st()))
        # the iteration after a reset iteration hands rejected rows old + new gradient (include/graspqp_b200.h,
        # gqp_mala_accept: after_reset)
        self._prev_reset, self._reset_now = self._reset_now, False
        with torch.no_grad():  # optimizer.py:325-338: the hand model now holds accepted-or-restored rows
            hm.hand_pose.data.copy_(pose)
            hm.contact_point_indices.copy_(idx.to(hm.contact_point_indices.dtype))
            if hm.hand_pose.grad is not None:
                hm.hand_pose.grad.copy_(grad)
            keep = hm.hand_pose.grad
            hm.set_parameters(hm.hand_pose.detach().requires_grad_(True), hm.contact_point_indices)
            hm.hand_pose.grad = keep
        self._noise = None
        return acc.bool(), T

    def reset_envs(self, mask):
        self.step[mask] = 0
        self.ema_grad_hand_pose[mask] = 0
        self._reset_now = True

    def zero_grad(self):
        if self.hand_model.hand_pose.grad is not None:
            self.hand_model.hand_pose.grad.data.zero_()


class AnnealingDexGraspNet(MalaStar):
    """optimizer.py:11-149 (--optimizer dexgraspnet): one RMS accumulator (D) and one step counter for the batch, no clipping,
    temperature without the z-score factor, reset_envs is a no-op and resets are not force-accepted."""

    def __init__(self, hand_model, switch_possibility=0.5, starting_temperature=18, temperature_decay=0.95,
                 annealing_period=30, step_size=0.005, stepsize_period=50, mu=0.98, device="cuda", seed=1, **kwargs):
        super().__init__(hand_model, switch_possibility, starting_temperature, temperature_decay, annealing_period,
                         step_size, stepsize_period, mu, device, clip_grad=False, seed=seed)
        D = hand_model.hand_pose.shape[1]
        self.step = 0
        self._ema2 = torch.zeros(2, D, device=hand_model.hand_pose.device)
        self._cur = 0

    @property
    def ema_grad_hand_pose(self):
        return self._ema2[self._cur] if hasattr(self, "_ema2") else None

    @ema_grad_hand_pose.setter
    def ema_grad_hand_pose(self, v):  # MalaStar.__init__ assigns its (B,D) accumulator: not used by this optimiser
        pass

    def try_step(self):
        lib, hm = _lib.lib(), self.hand_model
        pose = hm.hand_pose.detach().contiguous()
        grad = (hm.hand_pose.grad if hm.hand_pose.grad is not None else torch.zeros_like(pose)).contiguous()
        idx = hm.contact_point_indices.to(torch.int32).contiguous()
        B, D = pose.shape
        n_c = idx.shape[1]
        self.iteration += 1
        _lib.check(lib.gqp_mala_stats(grad.data_ptr(), self._z.data_ptr(), 1, B, D, 0, self._g2.data_ptr(),
                                      torch.empty_like(self._z).data_ptr(), self._scratch.data_ptr(), self._st()))
        pose_new, idx_new = torch.empty_like(pose), torch.empty_like(idx)
        s = torch.empty(1, device=pose.device)
        us, ir = (self._noise[0], self._noise[1]) if self._noise is not None else (None, None)
        _lib.check(lib.gqp_anneal_propose(ctypes.byref(self.params), pose.data_ptr(), grad.data_ptr(), self._g2.data_ptr(),
                                          self._ema2[self._cur].data_ptr(), self._ema2[1 - self._cur].data_ptr(), self.step,
                                          idx.data_ptr(), B, D, n_c, hm.n_contact_candidates,
                                          us.data_ptr() if us is not None else None,
                                          ir.data_ptr() if ir is not None else None, self.seed, self.iteration,
                                          pose_new.data_ptr(), idx_new.data_ptr(), s.data_ptr(), self._st()))
        self._cur = 1 - self._cur
        self.old_hand_pose, self.old_contact_point_indices = hm.hand_pose, hm.contact_point_indices
        self.old_grad_hand_pose = grad
        hm.set_parameters(pose_new.requires_grad_(True), idx_new.to(hm.contact_point_indices.dtype))
        self.step += 1
        return s[0]

    def accept_step(self, energy, new_energy, *args, **kwargs):
        lib, hm = _lib.lib(), self.hand_model
        B, D = hm.hand_pose.shape
        n_c = hm.contact_point_indices.shape[1]
        e_old = energy.detach().float().contiguous().clone()
        e_new = new_energy.detach().float().contiguous()
        ua = self._noise[2] if self._noise is not None else None
        acc = torch.empty(B, dtype=torch.uint8, device=e_new.device)
        T = torch.empty(B, device=e_new.device)
        pose = self.old_hand_pose.detach().clone().contiguous()
        idx = self.old_contact_point_indices.to(torch.int32).clone().contiguous()
        grad = self.old_grad_hand_pose.clone().contiguous()
        pose_new = hm.hand_pose.detach().contiguous()
        idx_new = hm.contact_point_indices.to(torch.int32).contiguous()
        grad_new = (hm.hand_pose.grad if hm.hand_pose.grad is not None else torch.zeros_like(pose_new)).contiguous()
        losses = torch.empty(1, B, device=e_new.device)
        _lib.check(lib.gqp_anneal_accept(ctypes.byref(self.params), B, D, n_c, 1, self._one.data_ptr(), e_new.data_ptr(),
                                         ua.data_ptr() if ua is not None else None, self.seed, self.iteration, self.step,
                                         pose_new.data_ptr(), idx_new.data_ptr(), grad_new.data_ptr(), pose.data_ptr(),
                                         idx.data_ptr(), grad.data_ptr(), e_old.data_ptr(), losses.data_ptr(), None,
                                         acc.data_ptr(), T.data_ptr(), int(self._prev_reset), self._st()))
        self._prev_reset, self._reset_now = self._reset_now, False
        with torch.no_grad():
            hm.hand_pose.data.copy_(pose)
            hm.contact_point_indices.copy_(idx.to(hm.contact_point_indices.dtype))
            if hm.hand_pose.grad is not None:
                hm.hand_pose.grad.copy_(grad)
            keep = hm.hand_pose.grad
            hm.set_parameters(hm.hand_pose.detach().requires_grad_(True), hm.contact_point_indices)
            hm.hand_pose.grad = keep
        self._noise = None
        return acc.bool(), T[0]

    def reset_envs(self, mask):  # :148-149 is `pass`; the iteration after a reset still sees the gradient rule
        self._reset_now = True
