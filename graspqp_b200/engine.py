"""FitEngine — the loop body of the reference's scripts/fit.py:399-458 as a sequence of C-ABI kernel launches.

One engine = one device = one group of objects (the reference's job-level sharding, scripts/fit_all.sh:154-248), with
`batch_each` grasp candidates per object.  All state lives in device buffers; an iteration enqueues ~14 launches on the
given stream and never synchronises, so it can be captured in a CUDA graph (see `FitEngine.capture`).

The engine is written against two tiny adapters so that the SAME host code drives
  * the product: libgraspqp_b200.so + torch CUDA tensors (`TorchBackend`), and
  * the tests:   the kernel sources run by the host SIMT emulator + numpy buffers (`NumpyBackend`, tests only).
There is no CPU implementation of any stage here: without one of those two libraries nothing computes.

Order of the energy terms (= the reference's dict order, core/energy.py:29-66): E_dis, E_fc, E_joints, E_pen, E_spen.
"""
import ctypes

import numpy as np

from . import _lib, hand_pack

TERMS = ("E_dis", "E_fc", "E_joints", "E_pen", "E_spen")
OPTIONAL_TERMS = ("E_prior", "E_wall")  # energy.py:68-77, evaluated when their weight is > 0 (fit.py:373), appended in this order
DEFAULT_WEIGHTS = dict(E_dis=100.0, E_fc=1.0, E_joints=1.0, E_pen=100.0, E_spen=10.0, E_prior=0.0,
                       E_wall=0.0)  # scripts/fit.py:77-78, 90-112


def morton_order(pts, bits=10):
    """permutation sorting (N,3) points along a Z-order curve"""
    p = np.asarray(pts, np.float64)
    lo, hi = p.min(0), p.max(0)
    q = np.minimum(((p - lo) / np.maximum(hi - lo, 1e-12) * (1 << bits)).astype(np.int64), (1 << bits) - 1)
    code = np.zeros(len(p), np.int64)
    for b in range(bits):
        for k in range(3):
            code |= ((q[:, k] >> b) & 1) << (3 * b + k)
    return np.argsort(code, kind="stable")


class NumpyBackend:
    """Host buffers for the emulator library (tests/_emu).  Test infrastructure."""
    stream = None

    def empty(self, shape, dtype):
        return np.zeros(shape, dtype)

    def from_host(self, a, dtype=None):
        return np.array(a, dtype=dtype, copy=True, order="C")  # an upload is a copy: the caller's array stays its own

    def to_host(self, a):
        return np.array(a, copy=True)

    def ptr(self, a):
        return None if a is None else a.ctypes.data

    def copy_(self, dst, src):
        dst[...] = src

    side_stream = None  # the emulator runs launches in order: no second stream

    def fork(self):
        pass

    def join(self):
        pass


class TorchBackend:
    """torch CUDA tensors; launches go to torch's current stream."""

    def __init__(self, device="cuda"):
        import torch
        self.torch = torch
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("graspqp_b200 computes on CUDA devices only (no CPU fallback)")

    _np2t = {np.float32: "float32", np.int32: "int32", np.uint8: "uint8", np.float64: "float64", np.int64: "int64"}

    @property
    def stream(self):
        return ctypes.c_void_p(self.torch.cuda.current_stream(self.device).cuda_stream)

    def empty(self, shape, dtype):
        return self.torch.zeros(shape, dtype=getattr(self.torch, self._np2t[dtype]), device=self.device)

    def from_host(self, a, dtype=None):
        t = self.torch.as_tensor(np.ascontiguousarray(a, dtype))
        return t.to(self.device).contiguous()

    def to_host(self, a):
        return a.detach().cpu().numpy()

    def ptr(self, a):
        return None if a is None else a.data_ptr()

    def copy_(self, dst, src):
        dst.copy_(src)

    # ---- second stream: the penetration chain runs beside the contact / force-closure chain ----
    def _side(self):
        if getattr(self, "_side_obj", None) is None:
            self._side_obj = self.torch.cuda.Stream(device=self.device)
            self._ev_fork = self.torch.cuda.Event()
            self._ev_join = self.torch.cuda.Event()
        return self._side_obj

    @property
    def side_stream(self):
        return ctypes.c_void_p(self._side().cuda_stream)

    def fork(self):
        """work enqueued on the side stream from now on sees everything enqueued on the current stream so far"""
        side = self._side()
        self._ev_fork.record(self.torch.cuda.current_stream(self.device))
        side.wait_event(self._ev_fork)

    def join(self):
        """the current stream waits for everything enqueued on the side stream"""
        self._ev_join.record(self._side())
        self.torch.cuda.current_stream(self.device).wait_event(self._ev_join)


class FitEngine:
    def __init__(self, cdll, backend, pack, face_verts_list, surface_points, batch_each, hand_pose, contact_idx,
                 weights=None, friction=0.2, max_limit=20.0, svd_gain=0.1, torque_weight=5.0, switch_possibility=0.4,
                 starting_temperature=18.0, temperature_decay=0.95, annealing_period=30, step_size=0.005,
                 stepsize_period=50, mu=0.98, clip_grad=False, qp_batch_stop=True, seed=1, collectives=None,
                 overlap=True, queue_per_candidate=None, energy_type="graspqp", optimizer="mala_star",
                 tdg_directions=None, dexgrasp_torque_weight=0.0, hand_surface=None):
        """energy_type: graspqp | dexgrasp | tdg (scripts/fit.py:81-88, 344-356); optimizer: mala_star | dexgraspnet
        (fit.py:74, 334-337).  tdg_directions (P,3): the unit directions TDGEnergy draws at construction (tdg.py:148-150);
        default: 1000 normalised normal draws from numpy's RandomState(seed), as np.random.seed(seed) makes the reference's.
        dexgrasp_torque_weight: 0.0 is what the reference's call passes (DexgraspSpanMetric.forward's default).
        weights E_prior / E_wall > 0 switch those terms on (energy.py:68-77); hand_surface = (points (n,3) in link frames,
        link (n)) for E_wall, default hand_pack.surface_points(pack) (512 samples, built the reference's way)."""
        if energy_type not in ("graspqp", "dexgrasp", "tdg"):
            raise ValueError(f"energy_type {energy_type!r}: graspqp, dexgrasp or tdg")
        if optimizer not in ("mala_star", "dexgraspnet"):
            raise ValueError(f"optimizer {optimizer!r}: mala_star or dexgraspnet")
        self.energy_type, self.optimizer = energy_type, optimizer
        self.lib, self.be = cdll, backend
        self.pack = pack
        self.n_obj = len(face_verts_list)
        self.batch_each = int(batch_each)
        self.Bt = self.n_obj * self.batch_each
        self.dof = len(pack["joints_lower"])
        self.D = 9 + self.dof
        self.L = len(pack["mesh_frame"])
        self.n_cand = len(pack["cand"])
        self.n_c = int(np.asarray(contact_idx).shape[1])
        self.seed = int(seed)
        self.iteration = 0
        self.qp_batch_stop = bool(qp_batch_stop)
        self.overlap = bool(overlap)
        self._graph = None
        self.graph_launches = 0
        self.collectives = collectives  # shard.GlobalStats or None (no data-path collective: the default)
        be, Bt, D, n_c, L = backend, self.Bt, self.D, self.n_c, self.L
        surface_points = np.ascontiguousarray(surface_points, np.float32)
        assert surface_points.shape[0] == self.n_obj and surface_points.shape[2] == 3
        self.n_pts = surface_points.shape[1]
        # spatial (Morton) order: the penetration kernel culls whole 32-sample chunks against each link's box
        surface_points_sorted = np.stack([sp[morton_order(sp)] for sp in surface_points])
        # ---- handles ----
        self.hand = hand_pack.create_hand(cdll, pack)
        self.meshes = []
        for fv in face_verts_list:
            fv = np.ascontiguousarray(fv, np.float32).reshape(-1, 9)
            m = ctypes.c_void_p()
            _lib.check(cdll.gqp_mesh_create(fv.ctypes.data, len(fv), ctypes.byref(m)), cdll)
            self.meshes.append(m)
        arr = (ctypes.c_void_p * self.n_obj)(*[m.value for m in self.meshes])
        self.objset = ctypes.c_void_p()
        _lib.check(cdll.gqp_objset_create(ctypes.cast(arr, ctypes.c_void_p), self.n_obj, ctypes.byref(self.objset)), cdll)
        # ---- parameters ----
        w = dict(DEFAULT_WEIGHTS)
        w.update(weights or {})
        self.weights = w
        self.fc = _lib.GqpFcParams(friction=friction, torque_weight=torque_weight, lower=1.0, upper=max_limit + 1.0,
                                   ridge=1e-4, svd_gain=svd_gain, values_gain=2.0, values_offset=1e-2, eps=5e-2,
                                   max_iter=12, w_fc=w["E_fc"], w_dis=w["E_dis"])
        self.mala = _lib.GqpMalaParams(switch_possibility=switch_possibility, starting_temperature=starting_temperature,
                                       temperature_decay=temperature_decay, step_size=step_size, mu=mu,
                                       annealing_period=annealing_period, stepsize_period=stepsize_period,
                                       clip_grad=int(clip_grad))
        # ---- device state ----
        f32, i32, u8 = np.float32, np.int32, np.uint8
        self.surf = be.from_host(surface_points_sorted, f32)
        self.cog = be.from_host(surface_points.mean(1, dtype=np.float32), f32)  # object_model.py:64-68
        self.terms = TERMS + tuple(k for k in OPTIONAL_TERMS if w[k] > 0.0)
        self.w_dev = be.from_host(np.array([w[k] for k in self.terms], f32))
        self.n_surf = 0
        if len(self.terms) > len(TERMS):
            ax = {"x": (1, 0, 0), "y": (0, 1, 0), "z": (0, 0, 1), "-x": (-1, 0, 0), "-y": (0, -1, 0), "-z": (0, 0, -1)}
            self.grasp_axis = (ctypes.c_float * 3)(*ax[str(pack["axes"][2])])  # HandModel.grasp_axis, hand_model.py:470
            self.hand_surf = self.hand_surf_link = None
            if "E_wall" in self.terms:
                sp_, sl_ = hand_surface if hand_surface is not None else hand_pack.surface_points(pack)
                assert np.asarray(sl_).min() >= 0 and np.asarray(sl_).max() < self.L
                self.hand_surf, self.hand_surf_link = be.from_host(sp_, f32), be.from_host(sl_, i32)
                self.n_surf = len(sl_)
        self.pose = be.from_host(hand_pose, f32)
        self.idx = be.from_host(contact_idx, i32)
        assert tuple(self.pose.shape) == (Bt, D) and tuple(self.idx.shape) == (Bt, n_c)
        self.grad, self.ema = be.empty((Bt, D), f32), be.empty((Bt, D), f32)
        self.step_count = be.empty((Bt,), i32)
        self.energy, self.losses = be.empty((Bt,), f32), be.empty((len(self.terms), Bt), f32)
        self.pose_new, self.idx_new = be.empty((Bt, D), f32), be.empty((Bt, n_c), i32)
        self.grad_new, self.losses_new = be.empty((Bt, D), f32), be.empty((len(self.terms), Bt), f32)
        self.new_energy, self.accept = be.empty((Bt,), f32), be.empty((Bt,), u8)
        self.temperature, self.step_size_out = be.empty((Bt,), f32), be.empty((Bt,), f32)
        self.z, self.grad2_mean = be.empty((Bt,), f32), be.empty((D,), f32)
        self.stats_scratch = be.empty(((Bt + 63) // 64 * D,), f32)
        self.link_tf, self.g_link_tf = be.empty((Bt, L, 12), f32), be.empty((Bt, L, 12), f32)
        self.cpts, self.cnrm = be.empty((Bt, n_c, 3), f32), be.empty((Bt, n_c, 3), f32)
        self.g_cpts, self.g_cnrm = be.empty((Bt, n_c, 3), f32), be.empty((Bt, n_c, 3), f32)
        self.obj_dist, self.obj_nrm = be.empty((Bt, n_c), f32), be.empty((Bt, n_c, 3), f32)
        self.obj_clst, self.obj_sign = be.empty((Bt, n_c, 3), f32), be.empty((Bt, n_c), i32)
        self.obj_dgrad = be.empty((Bt, n_c, 3), f32)
        self.ones = be.from_host(np.ones((Bt, n_c), f32))
        self.lam = be.empty((Bt, n_c), f32)
        self.g_pose_joints = be.empty((Bt, D), f32)
        # survivor queue of the penetration stage: unbounded (n_pts per candidate, 20 KB each) for small batches, 512 per
        # candidate on average from 64k candidates on (20-45 are observed along a fit; overflow is detected, assert_ok)
        self.queue_per_candidate = int(queue_per_candidate if queue_per_candidate is not None else
                                       (0 if Bt < 65536 else 512))
        self.pen_scratch = be.empty((int(cdll.gqp_hand_distance_scratch_floats_bounded(
            self.n_obj, self.batch_each, self.n_pts, 0, self.queue_per_candidate)),), f32)
        self.pen_counter_off = int(cdll.gqp_hand_distance_counter_offset(self.n_obj, self.batch_each, self.n_pts, 0,
                                                                         be.ptr(self.pen_scratch)))
        self.qp_stop = be.empty((1,), i32)
        if energy_type == "graspqp":
            self.qp_ws = be.empty(((int(cdll.gqp_fc_workspace_bytes(Bt, n_c, self.fc.max_iter)) + 7) // 8,), np.float64)
        else:
            self.span = _lib.GqpSpanParams(kind=_lib.SPAN_DEXGRASP if energy_type == "dexgrasp" else _lib.SPAN_TDG,
                                           torque_weight=dexgrasp_torque_weight, friction=0.2, obb_length=0.2, scale=100.0,
                                           enable_density=1, w_fc=w["E_fc"], w_dis=w["E_dis"])  # tdg.py:219-227, :236
            self.tdg_dirs, self.n_dirs = None, 0
            if energy_type == "tdg":
                if tdg_directions is None:  # tdg.py:44-48 after np.random.seed(seed) (fit.py:159)
                    d = np.random.RandomState(self.seed).randn(1000, 3)
                    tdg_directions = d / (np.linalg.norm(d, axis=-1)[:, None] + 1e-8)
                tdg_directions = np.ascontiguousarray(tdg_directions, f32).reshape(-1, 3)
                self.tdg_dirs, self.n_dirs = be.from_host(tdg_directions, f32), len(tdg_directions)
        if optimizer == "dexgraspnet":  # one RMS accumulator for the batch, double-buffered; one step counter (host)
            self.ema = be.empty((2, D), f32)
            self.ema_cur, self.anneal_step = 0, 0
            self.mala.clip_grad = 0  # AnnealingDexGraspNet has none (optimizer.py:70-108)
        self.prev_reset = False
        self.reset_all = be.from_host(np.ones(Bt, u8))
        self.reset_mask, self.reset_count = be.empty((Bt,), u8), be.empty((1,), i32)
        self.idx_eval = be.empty((Bt, n_c), i32)
        self._init_energy()

    # ------------------------------------------------------------------------------------------------------------
    def _ck(self, rc):
        _lib.check(rc, self.lib)

    def evaluate(self, pose, idx, losses_out, grad_out):
        """calculate_energy (core/energy.py:6-66) + backward (fit.py:440) at (pose, idx): 5 terms and d(sum w E)/d pose."""
        lib, be, p, st = self.lib, self.be, self.be.ptr, self.be.stream
        Bt, n_c, O, B = self.Bt, self.n_c, self.n_obj, self.batch_each
        row = lambda k: p(losses_out) + 4 * Bt * self.terms.index(k) if k in self.terms else None
        # H1-H3 in one launch, a warp per candidate: link transforms + the n_c chosen contact points / normals
        self._ck(lib.gqp_fk_contacts_forward(self.hand, p(pose), p(idx), Bt, n_c, p(self.link_tf), p(self.cpts),
                                             p(self.cnrm), st))
        self._ck(lib.gqp_memset_zero(p(self.g_link_tf), 4 * Bt * self.L * 12, st))
        # two independent chains read link_tf: the penetration chain (which adds into g_link_tf with atomics) goes to a
        # second stream so that its latency-bound kernels fill the SMs beside the contact / force-closure chain
        overlap = self.overlap and be.side_stream is not None
        s2 = be.side_stream if overlap else st
        if overlap:
            be.fork()
        self._ck(lib.gqp_object_distance_forward(self.objset, B, n_c, p(self.cpts), 1.0, p(self.obj_dist),
                                                 p(self.obj_nrm), p(self.obj_clst), p(self.obj_sign), st))
        self._ck(lib.gqp_object_distance_backward(p(self.ones), p(self.cpts), p(self.obj_clst), p(self.obj_sign),
                                                  Bt * n_c, 1.0, p(self.obj_dgrad), st))
        if self.energy_type != "graspqp":  # (f3) dexgrasp / tdg energy + E_dis, one launch
            self._ck(lib.gqp_span_energy(ctypes.byref(self.span), p(self.cpts), p(self.cnrm), p(self.obj_dist),
                                         p(self.obj_nrm), p(self.obj_dgrad), p(self.cog), p(self.tdg_dirs), self.n_dirs, O,
                                         B, n_c, row("E_fc"), row("E_dis"), p(self.g_cpts), p(self.g_cnrm), st))
        else:
            self._evaluate_graspqp(row)
        self._evaluate_tail(pose, idx, grad_out, row, overlap, s2)

    def _evaluate_graspqp(self, row):
        lib, p, st = self.lib, self.be.ptr, self.be.stream
        n_c, O, B = self.n_c, self.n_obj, self.batch_each
        # Q1-Q4 + E_dis: one interior-point pass that records the batch statistics of qpth's exit test, then the
        # finish pass (value, svd term, energies, analytic backward) from every candidate's best iterate as of the exit
        self._ck(lib.gqp_fc_solve(ctypes.byref(self.fc), p(self.cpts), p(self.obj_nrm), p(self.cog), O, B, n_c,
                                  p(self.qp_ws), st))
        if self.collectives is not None and self.qp_batch_stop:  # qpth's exit test over the candidates of ALL ranks
            hdr = self.qp_ws.view(np.int32)[:128] if isinstance(self.qp_ws, np.ndarray) else \
                self.qp_ws.view(self.be.torch.int32)[:128]
            self.collectives.reduce_qp_header(hdr)
        self._ck(lib.gqp_fc_finish(ctypes.byref(self.fc), p(self.cpts), p(self.cnrm), p(self.obj_dist), p(self.obj_nrm),
                                   p(self.obj_dgrad), p(self.cog), O, B, n_c, p(self.qp_ws), int(self.qp_batch_stop),
                                   row("E_fc"), row("E_dis"), p(self.lam), p(self.g_cpts), p(self.g_cnrm),
                                   p(self.qp_stop), st))

    def _evaluate_tail(self, pose, idx, grad_out, row, overlap, s2):
        lib, be, p, st = self.lib, self.be, self.be.ptr, self.be.stream
        Bt, n_c, O, B = self.Bt, self.n_c, self.n_obj, self.batch_each
        # the side chain is enqueued AFTER the critical chain: its persistent CTAs then fill what the one-wave
        # object-distance / QP kernels leave free instead of holding the SMs while the critical chain waits
        self._ck(lib.gqp_hand_distance_bounded(self.hand, p(self.link_tf), p(self.surf), O, B, self.n_pts, 0,
                                               row("E_pen"), p(self.pen_scratch), self.weights["E_pen"],
                                               p(self.g_link_tf), self.queue_per_candidate, s2))
        self._ck(lib.gqp_self_pen_joints(self.hand, p(self.link_tf), p(pose), Bt, row("E_spen"), row("E_joints"),
                                         self.weights["E_spen"], self.weights["E_joints"], p(self.g_link_tf),
                                         p(self.g_pose_joints), s2))
        if len(self.terms) > len(TERMS):  # (f3) E_prior / E_wall: adds to g_link_tf and to g_pose_joints' rotation columns
            self._ck(lib.gqp_prior_wall(self.hand, p(self.link_tf), p(pose), Bt, ctypes.byref(self.grasp_axis),
                                        p(self.hand_surf), p(self.hand_surf_link), self.n_surf,
                                        self.weights["E_prior"], self.weights["E_wall"], row("E_prior"), row("E_wall"),
                                        p(self.g_link_tf), p(self.g_pose_joints), s2))
        if overlap:
            be.join()
        # contact scatter + analytic J' of the chain in one launch (fixed summation order): reads what the penetration
        # chain accumulated in g_link_tf, adds the contact terms, writes d(sum w E)/d pose
        self._ck(lib.gqp_fk_contacts_backward(self.hand, p(pose), p(idx), p(self.g_cpts), p(self.g_cnrm),
                                              p(self.g_link_tf), p(self.g_pose_joints), Bt, n_c, p(grad_out), st))

    def capture(self):
        """Record the energy evaluation of an iteration (both streams, ~20 launches) into a CUDA graph; step() then
        replays it.  The MALA* kernels stay ordinary launches (their iteration counter is a by-value argument).
        Not available with the emulator backend or with per-iteration collectives."""
        if self.collectives is not None or not hasattr(self.be, "torch"):
            return False
        torch = self.be.torch
        torch.cuda.synchronize(self.be.device)
        g = torch.cuda.CUDAGraph()
        cap = torch.cuda.Stream(device=self.be.device)
        cap.wait_stream(torch.cuda.current_stream(self.be.device))
        with torch.cuda.stream(cap):
            n0 = int(self.lib.gqp_launch_count())
            self.evaluate(self.pose_new, self.idx_new, self.losses_new, self.grad_new)  # warm-up outside the capture
            self.graph_launches = int(self.lib.gqp_launch_count()) - n0  # kernels one replay executes
            cap.synchronize()
            with torch.cuda.graph(g, stream=cap):
                self.evaluate(self.pose_new, self.idx_new, self.losses_new, self.grad_new)
        torch.cuda.current_stream(self.be.device).wait_stream(cap)
        self._graph = g
        return True

    def _accept(self, z, reset, u_accept, after_reset=False):
        lib, p, st = self.lib, self.be.ptr, self.be.stream
        if self.optimizer == "dexgraspnet":  # no z-score factor, no forced accept (optimizer.py:110-138)
            self._ck(lib.gqp_anneal_accept(ctypes.byref(self.mala), self.Bt, self.D, self.n_c, len(self.terms), p(self.w_dev),
                                           p(self.losses_new), p(u_accept), self.seed, self.iteration, self.anneal_step,
                                           p(self.pose_new), p(self.idx_new), p(self.grad_new), p(self.pose), p(self.idx),
                                           p(self.grad), p(self.energy), p(self.losses), p(self.new_energy),
                                           p(self.accept), p(self.temperature), int(after_reset), st))
            return
        self._ck(lib.gqp_mala_accept(ctypes.byref(self.mala), self.Bt, self.D, self.n_c, len(self.terms), p(self.w_dev),
                                     p(self.losses_new), p(z), p(reset), p(u_accept), self.seed, self.iteration,
                                     p(self.step_count), p(self.pose_new), p(self.idx_new), p(self.grad_new),
                                     p(self.pose), p(self.idx), p(self.grad), p(self.energy), p(self.losses),
                                     p(self.new_energy), p(self.accept), p(self.temperature), int(after_reset), st))

    def _init_energy(self):
        """fit.py:381-396: first energy + backward, then optimizer.zero_grad()."""
        self.be.copy_(self.pose_new, self.pose)
        self.be.copy_(self.idx_new, self.idx)
        self.evaluate(self.pose_new, self.idx_new, self.losses_new, self.grad_new)
        if self.optimizer == "dexgraspnet":  # commit everything: u = -1 < exp(.) always
            self._accept(None, None, self.be.from_host(np.full(self.Bt, -1.0, np.float32)))
        else:
            self._accept(None, self.reset_all, None)
        self.first_grad = self.be.to_host(self.grad)
        self._ck(self.lib.gqp_memset_zero(self.be.ptr(self.grad), 4 * self.Bt * self.D, self.be.stream))

    def step(self, u_switch=None, idx_rand=None, u_accept=None, reset_init=None, reset_mask=None,
             z_score_threshold=1.0):
        """One iteration of fit.py:399-458.  Noise arrays (device buffers) are injected, or drawn by Philox if None.

        reset_init = (pose_init (Bt,D), idx_init (Bt,n_c)) makes this a RESET iteration (fit.py:408-422): rows with
        z-score > z_score_threshold (or the injected uint8 `reset_mask`) are re-initialised from it between the proposal
        and the energy evaluation, their step counter and RMS accumulator restart (MalaStar.reset_envs) and their
        proposal is accepted unconditionally.  As in the reference (hand_model.py:830-831), the energies of a reset
        iteration are evaluated at the freshly drawn contact indices for EVERY row."""
        lib, p, st = self.lib, self.be.ptr, self.be.stream
        self.iteration += 1
        after_reset, self.prev_reset = self.prev_reset, reset_init is not None
        self._ck(lib.gqp_mala_stats(p(self.grad), p(self.energy), self.n_obj, self.batch_each, self.D,
                                    self.mala.clip_grad, p(self.grad2_mean), p(self.z), p(self.stats_scratch), st))
        if self.collectives is not None:  # (g**2).mean(0) over the rows of ALL ranks (optimizer.py:231)
            self.collectives.reduce_rms(self.grad2_mean, self.Bt)
        if self.optimizer == "dexgraspnet":  # AnnealingDexGraspNet.try_step (optimizer.py:70-108)
            row = lambda k: p(self.ema) + 4 * self.D * k
            self._ck(lib.gqp_anneal_propose(ctypes.byref(self.mala), p(self.pose), p(self.grad), p(self.grad2_mean),
                                            row(self.ema_cur), row(1 - self.ema_cur), self.anneal_step, p(self.idx),
                                            self.Bt, self.D, self.n_c, self.n_cand, p(u_switch), p(idx_rand), self.seed,
                                            self.iteration, p(self.pose_new), p(self.idx_new), p(self.step_size_out),
                                            st))
            self.ema_cur, self.anneal_step = 1 - self.ema_cur, self.anneal_step + 1
        else:
            self._ck(lib.gqp_mala_propose(ctypes.byref(self.mala), p(self.pose), p(self.grad), p(self.grad2_mean),
                                          p(self.ema), p(self.step_count), p(self.idx), self.Bt, self.D, self.n_c,
                                          self.n_cand, p(u_switch), p(idx_rand), self.seed, self.iteration,
                                          p(self.pose_new), p(self.idx_new), p(self.step_size_out), st))
        if reset_init is not None:
            pose_init, idx_init = reset_init
            ema = self.ema
            if self.optimizer == "dexgraspnet":  # reset_envs is a no-op there (:148-149): the kernel's per-row restart
                if getattr(self, "_reset_sink", None) is None:  # of step / ema goes to a sink
                    self._reset_sink = self.be.empty((self.Bt, self.D), np.float32)
                ema = self._reset_sink
            self._ck(lib.gqp_mala_reset(self.Bt, self.D, self.n_c, p(self.z), float(z_score_threshold), p(reset_mask),
                                        p(pose_init), p(idx_init), p(self.pose_new), p(self.idx_new), p(self.idx_eval),
                                        p(self.step_count), p(ema), p(self.reset_mask), p(self.reset_count), st))
            self.evaluate(self.pose_new, self.idx_eval, self.losses_new, self.grad_new)
            self._accept(self.z, self.reset_mask, u_accept, after_reset)
            return
        if self._graph is not None:
            self._graph.replay()
        else:
            self.evaluate(self.pose_new, self.idx_new, self.losses_new, self.grad_new)
        self._accept(self.z, None, u_accept, after_reset)

    # ---- (f1) outlier resets with the re-initialisation on the device --------------------------------------------------
    def enable_device_init(self, hull_tris_list, **params):
        """hull_tris_list: outward-oriented convex-hull triangles per object (initializations.convex_hull)."""
        from . import initializations
        self.init = initializations.DeviceInit(self.lib, self.be, self.pack, hull_tris_list, self.batch_each, self.n_c,
                                               **params)
        self.pose_init = self.be.empty((self.Bt, self.D), np.float32)
        self.idx_init = self.be.empty((self.Bt, self.n_c), np.int32)

    def reset_step(self, z_score_threshold=1.0, u_pool=None, u_pose=None, idx_rand_init=None, **step_noise):
        """A reset iteration (fit.py:408-422): initialize_convex_hull for every row on the device, then step() with it."""
        self.init.run(self.hand, self.pose_init, self.idx_init, u_pool, u_pose, idx_rand_init, seed=self.seed)
        self.step(reset_init=(self.pose_init, self.idx_init), z_score_threshold=z_score_threshold, **step_noise)

    def fit(self, n_iter, reset_epochs=600, z_score_threshold=1.0):
        """scripts/fit.py:399-458 for n_iter iterations, counted from this engine's first one: resets every reset_epochs
        iterations except in the last 2*reset_epochs (fit.py:408)."""
        for step in range(self.iteration + 1, n_iter + 1):
            if reset_epochs and getattr(self, "init", None) is not None and step % reset_epochs == 0 \
                    and step < n_iter - 2 * reset_epochs:
                self.reset_step(z_score_threshold)
            else:
                self.step()

    def entropy(self):
        """(f4) HandModel.joint_entropy / pose_entropy of the current batch (hand_model.py:698-760; fit.py:462-471 logs them
        with --log_entropy) -> dict with joints / translation / rotation entropy and fit.py's `total`; synchronises."""
        be, p = self.be, self.be.ptr
        if getattr(self, "_ent", None) is None:
            self._ent = (be.empty((self.dof + 6,), np.float32), be.empty((3,), np.float32),
                         be.from_host(self.pack["joints_lower"], np.float32),
                         be.from_host(self.pack["joints_upper"], np.float32))
        scratch, out, lo, hi = self._ent
        self._ck(self.lib.gqp_entropy_stats(p(self.pose), self.Bt, self.dof, p(lo), p(hi), p(scratch), p(out), be.stream))
        j, t, r = [float(v) for v in be.to_host(out)]
        return {"joints_entropy": j, "translation_entropy": t, "rotation_entropy": r, "total": 0.5 * j + 0.5 * (t + r)}

    def pen_counters(self):
        """(samples queued for the exact search, work items, overflow flag) of the last penetration stage; synchronises"""
        v = self.be.to_host(self.pen_scratch.view(np.int32) if isinstance(self.pen_scratch, np.ndarray)
                            else self.pen_scratch.view(self.be.torch.int32)[self.pen_counter_off:self.pen_counter_off + 3])
        if isinstance(self.pen_scratch, np.ndarray):
            v = v[self.pen_counter_off:self.pen_counter_off + 3]
        return int(v[0]), int(v[1]), int(v[2])

    def assert_ok(self):
        """Raises if the bounded survivor queue of the last evaluated iteration overflowed (its E_pen is invalid)."""
        q, _, over = self.pen_counters()
        if over:
            raise _lib.GqpError(f"penetration survivor queue overflowed ({q} samples queued, capacity "
                                f"{self.queue_per_candidate} per candidate): construct FitEngine with a larger "
                                "queue_per_candidate (0 = unbounded)")

    _STATE = ("pose", "idx", "grad", "ema", "step_count", "energy", "losses")

    def snapshot(self):
        """copy of the optimiser state scripts/fit.py carries between iterations (pose, contact indices, gradient, RMS
        accumulator, step counters, energies) + the iteration number; see restore()"""
        extra = {"iteration": self.iteration, "prev_reset": self.prev_reset}
        if self.optimizer == "dexgraspnet":
            extra.update(ema_cur=self.ema_cur, anneal_step=self.anneal_step)
        return {k: self.be.to_host(getattr(self, k)) for k in self._STATE} | extra

    def restore(self, snap):
        for k in self._STATE:
            self.be.copy_(getattr(self, k), self.be.from_host(snap[k], snap[k].dtype))
        self.iteration = int(snap["iteration"])
        self.prev_reset = bool(snap.get("prev_reset", False))
        if self.optimizer == "dexgraspnet":
            self.ema_cur, self.anneal_step = int(snap["ema_cur"]), int(snap["anneal_step"])

    def stage_calls(self):
        """name -> callable re-running ONE stage of the iteration on the current buffers (per-stage timing in bench.py
        and tools/bench_kernels.py; results are overwritten with the same values, gradients accumulate garbage)."""
        lib, p, st = self.lib, self.be.ptr, self.be.stream
        Bt, n_c, O, B = self.Bt, self.n_c, self.n_obj, self.batch_each
        row = lambda k: p(self.losses_new) + 4 * Bt * TERMS.index(k)
        calls = {
            "fk_contacts_forward": lambda: lib.gqp_fk_contacts_forward(
                self.hand, p(self.pose_new), p(self.idx_new), Bt, n_c, p(self.link_tf), p(self.cpts), p(self.cnrm), st),
            "object_distance": lambda: lib.gqp_object_distance_forward(
                self.objset, B, n_c, p(self.cpts), 1.0, p(self.obj_dist), p(self.obj_nrm), p(self.obj_clst),
                p(self.obj_sign), st),
            "fc_solve": lambda: lib.gqp_fc_solve(ctypes.byref(self.fc), p(self.cpts), p(self.obj_nrm), p(self.cog), O, B,
                                                 n_c, p(self.qp_ws), st),
            "fc_finish": lambda: lib.gqp_fc_finish(
                ctypes.byref(self.fc), p(self.cpts), p(self.cnrm), p(self.obj_dist), p(self.obj_nrm), p(self.obj_dgrad),
                p(self.cog), O, B, n_c, p(self.qp_ws), int(self.qp_batch_stop), row("E_fc"), row("E_dis"), p(self.lam),
                p(self.g_cpts), p(self.g_cnrm), p(self.qp_stop), st),
            "hand_distance": lambda: lib.gqp_hand_distance_bounded(
                self.hand, p(self.link_tf), p(self.surf), O, B, self.n_pts, 0, row("E_pen"), p(self.pen_scratch),
                self.weights["E_pen"], p(self.g_link_tf), self.queue_per_candidate, st),
            "self_pen_joints": lambda: lib.gqp_self_pen_joints(
                self.hand, p(self.link_tf), p(self.pose_new), Bt, row("E_spen"), row("E_joints"), self.weights["E_spen"],
                self.weights["E_joints"], p(self.g_link_tf), p(self.g_pose_joints), st),
            "fk_contacts_backward": lambda: lib.gqp_fk_contacts_backward(
                self.hand, p(self.pose_new), p(self.idx_new), p(self.g_cpts), p(self.g_cnrm), p(self.g_link_tf),
                p(self.g_pose_joints), Bt, n_c, p(self.grad_new), st),
        }
        if self.energy_type != "graspqp":
            del calls["fc_solve"], calls["fc_finish"]
            calls["span_energy"] = lambda: lib.gqp_span_energy(
                ctypes.byref(self.span), p(self.cpts), p(self.cnrm), p(self.obj_dist), p(self.obj_nrm), p(self.obj_dgrad),
                p(self.cog), p(self.tdg_dirs), self.n_dirs, O, B, n_c, row("E_fc"), row("E_dis"), p(self.g_cpts),
                p(self.g_cnrm), st)
        return calls

    def close(self):
        lib = self.lib
        if getattr(self, "init", None) is not None:
            self.init.close()
            self.init = None
        if getattr(self, "objset", None):
            lib.gqp_objset_destroy(self.objset)
            self.objset = None
        for m in getattr(self, "meshes", []):
            lib.gqp_mesh_destroy(m)
        self.meshes = []
        if getattr(self, "hand", None):
            lib.gqp_hand_destroy(self.hand)
            self.hand = None
