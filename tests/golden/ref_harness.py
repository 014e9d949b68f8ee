"""TEST INFRASTRUCTURE — runs ONLY in the build container, where /root/reference exists.

Loads the reference's OWN Python modules for the hot path from where they lie and executes them on the CPU:

    graspqp/src/graspqp/core/optimizer.py      (MalaStar.try_step / accept_step / zero_grad)
    graspqp/src/graspqp/core/energy.py         (calculate_energy)
    graspqp/src/graspqp/metrics/ops/span.py    (friction cone, grasp matrix, OverallFrictionConeSpanMetric)
    graspqp/src/graspqp/metrics/ops/registry.py (SpanMetricWrapper, GraspSpanMetricFactory)
    graspqp/src/graspqp/metrics/solver/qp_solver.py (SQPLsqSolver: QP construction around qpth)
    graspqp/src/graspqp/utils/transforms.py    (robust_compute_rotation_matrix_from_ortho6d)
    thirdparty/pytorch_kinematics/pytorch_kinematics/{frame,chain,jacobian}.py + transforms/ (FK, Transform3d)

The package __init__ files are bypassed (they import trimesh / pytorch3d / lxml, absent here) by pre-registering empty
package modules whose __path__ points at the reference directories.  Third-party modules that are NOT vendored in the
reference and not installed are replaced by stubs built from oracle/:

    qpth.qp.QPFunction        -> oracle.qp.QPFunction  (restatement of qpth 0.0.18's batched PDIPM; PARITY UNPINNED)
    roma.special_gramschmidt  -> oracle.hand.special_gramschmidt (restatement of roma 1.5.4;        PARITY UNPINNED)

HandModel / ObjectModel themselves cannot be constructed here (trimesh, pytorch3d); the reference optimizer and energy
code are therefore driven with oracle.hand.HandOracle / oracle.pipeline.ObjectOracle as duck-typed models.  What the
golden vectors produced through this harness pin is the reference's own glue: MALA* update and accept rules, energy
assembly, friction cone, grasp matrix, QP construction, E_fc wrapper, pk forward kinematics and transforms.
"""
import os
import sys
import types

import numpy as np

import torch

REF = os.environ.get("GRASPQP_REFERENCE", "/root/reference")
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def available():
    return os.path.isdir(os.path.join(REF, "graspqp", "src", "graspqp"))


def load():
    """-> dict of reference modules (optimizer, energy, span, registry, qp_solver, transforms, pk_chain, pk_frame, pk_tf)"""
    if not available():
        raise RuntimeError("reference not present at " + REF)
    sys.path.insert(0, os.path.join(ROOT, "baseline"))
    import ref_loader
    return ref_loader.load(os.path.join(REF, "graspqp", "src", "graspqp"),
                           os.path.join(REF, "thirdparty", "pytorch_kinematics", "pytorch_kinematics"))


class _Noise:
    """Replaces torch inside graspqp.core.optimizer: rand / randint return the injected draws (optimizer.py:252-257,
    :311), everything else is the real torch."""

    def __init__(self):
        self.u_switch = self.idx_new = self.u_accept = self.mask_for_randint = None

    def __getattr__(self, k):
        return getattr(torch, k)

    def rand(self, *shape, **kw):
        if len(shape) == 2:
            return self.u_switch.to(kw.get("dtype", torch.float))
        return self.u_accept.to(kw.get("dtype", torch.float))

    def randint(self, high, size, **kw):
        # the reference draws exactly switch_mask.sum() values and scatters them in row-major mask order; the injected
        # (B,n_c) table is read in that same order
        assert self.idx_new.max() < high
        return self.idx_new[self.mask_for_randint]


def tdg_metric(ref, dtype, seed=7):
    """the reference's TDGSpanMetric with numpy seeded as scripts/fit.py:159 does before it draws its 1000 directions; for an
    fp64 run the fp32 constants TDGEnergy built (tdg.py:66-81, 148-153) are cast, nothing else is touched"""
    import numpy as np
    np.random.seed(seed)
    m = ref["registry"].TDGSpanMetric()
    te = m.tdg_energy
    if dtype != torch.float32:
        for k in ("rot_base1", "rot_base2", "miu_tensor", "target_direction_6D", "F_center_direction",
                  "obj_gravity_center", "obj_obb_length"):
            setattr(te, k, getattr(te, k).to(dtype))
    return m


def make_fit_state(pack, face_verts_list, surface_points_list, batch_each, hand_pose0, idx0, dtype=torch.float32,
                   friction=0.2, max_limit=20.0, n_cone_vecs=4, svd_gain=0.1, energy_type="graspqp",
                   optimizer="mala_star", weights=None, hand_surface=None, models="oracle", **optim):
    """The state scripts/fit.py holds between iterations, with the REFERENCE MalaStar / calculate_energy / metric."""
    from oracle.hand import HandOracle
    from oracle.pipeline import ObjectOracle, WEIGHTS
    ref = load()
    if models == "reference":
        # the reference's OWN HandModel / ObjectModel (instances without constructors, filled from the pack): every method
        # the loop calls — set_parameters, fk, cal_distance, self_penetration, ObjectModel.cal_distance — is the reference's
        assert hand_surface is None
        hand = reference_hand_model(pack, dtype)
        obj = reference_object_model(face_verts_list, surface_points_list, batch_each, dtype)
    else:
        hand = HandOracle(pack, dtype)
        if hand_surface is not None:
            hand.set_surface_points(*hand_surface)
        obj = ObjectOracle(face_verts_list, surface_points_list, batch_each, dtype)
        obj.device = torch.device("cpu")
        obj.object_scale_tensor = torch.ones(len(face_verts_list), batch_each, dtype=dtype)
    hand.set_parameters(torch.as_tensor(hand_pose0, dtype=dtype).clone().requires_grad_(True),
                        torch.as_tensor(idx0).long())
    cfg = dict(switch_possibility=0.4, starting_temperature=18.0, temperature_decay=0.95, annealing_period=30,
               step_size=0.005, stepsize_period=50, mu=0.98)  # scripts/fit.py argparse defaults
    cfg.update(optim)
    if optimizer == "mala_star":
        opt = ref["optimizer"].MalaStar(hand, device="cpu", batch_size=batch_each, **cfg)
    else:  # scripts/fit.py:336-337
        opt = ref["optimizer"].AnnealingDexGraspNet(hand, device="cpu", batch_size=batch_each, **cfg)
        if dtype != torch.float32:  # the class builds its scalars and accumulator as torch.float (optimizer.py:52-68)
            for k in ("starting_temperature", "temperature_decay", "step_size", "mu", "ema_grad_hand_pose"):
                setattr(opt, k, getattr(opt, k).to(dtype))
    F = ref["registry"].GraspSpanMetricFactory
    if energy_type == "graspqp":  # scripts/fit.py:344-356
        energy_fnc = F.create(F.MetricType.GRASPQP, solver_kwargs={"friction": friction, "max_limit": max_limit,
                                                                    "n_cone_vecs": n_cone_vecs})
    elif energy_type == "dexgrasp":
        energy_fnc = F.create(F.MetricType.DEXGRASP)
    elif energy_type == "tdg":
        energy_fnc = tdg_metric(ref, dtype)
    else:
        raise ValueError(energy_type)
    st = types.SimpleNamespace(hand=hand, obj=obj, opt=opt, energy_fnc=energy_fnc, weights=dict(WEIGHTS, **(weights or {})), B=batch_each,
                               svd_gain=svd_gain, ref=ref)

    def energies():
        losses = ref["energy"].calculate_energy(hand, obj, energy_fnc=energy_fnc,
                                                energy_names=[k for k, v in st.weights.items() if v > 0],
                                                method="gendexgrasp", svd_gain=svd_gain)
        e = 0
        for k, v in losses.items():
            e = e + st.weights[k] * v
        return losses, e

    st.energies = energies
    st.losses, st.energy = energies()          # fit.py:381-396
    st.energy.sum().backward()
    st.first_grad = hand.hand_pose.grad.clone()
    opt.zero_grad()
    return st


def fit_iterate(st, u_switch, idx_new, u_accept, reset=None, z_score_threshold=1.0):
    """One pass of scripts/fit.py:399-458 with the reference optimizer and energy code and injected noise.
    reset = (pose_init, idx_init, mask or None): a reset iteration (fit.py:408-422) — the reference's own
    MalaStar.reset_envs, with the result of initialize_convex_hull (trimesh / pytorch3d: absent here) injected through
    HandOracle.set_parameters(env_mask=...), the restatement of hand_model.py:845-850, 816-831."""
    ref, opt, hand = st.ref, st.opt, st.hand
    noise = _Noise()
    noise.u_switch, noise.idx_new, noise.u_accept = u_switch, idx_new, u_accept
    noise.mask_for_randint = u_switch < opt.switch_possibility
    mod = ref["optimizer"]
    real = mod.torch
    mod.torch = noise
    try:
        s = opt.try_step()
        Eb = st.energy.view(-1, st.B)
        z = ((Eb - Eb.mean(-1).unsqueeze(-1)) / Eb.std(-1).unsqueeze(-1)).view(-1)
        reset_mask = None
        if reset is not None:
            pose_init, idx_init, mask = reset
            reset_mask = (z > z_score_threshold) if mask is None else mask.bool()
            hand.set_parameters(pose_init.to(hand.dtype).clone().requires_grad_(True), idx_init.long(), env_mask=reset_mask)
            opt.reset_envs(reset_mask)
        opt.zero_grad()
        new_losses, new_energy = st.energies()
        new_energy.sum().backward()
        out = dict(s=s.clone(), z=z.detach().clone(), proposal=hand.hand_pose.detach().clone(),
                   idx_prop=hand.contact_point_indices.clone(), new_energy=new_energy.detach().clone(),
                   new_grad=hand.hand_pose.grad.clone(),
                   new_losses={k: v.detach().clone() for k, v in new_losses.items()})
        with torch.no_grad():
            accept, t = opt.accept_step(st.energy, new_energy, reset_mask, z, z_score_threshold)
            st.energy[accept] = new_energy[accept]
            for k, v in new_losses.items():
                st.losses[k][accept] = v[accept]
        out.update(accept=accept.clone(), temperature=t.clone(), energy=st.energy.detach().clone(),
                   pose=hand.hand_pose.detach().clone(), idx=hand.contact_point_indices.clone(),
                   step=opt.step.clone() if torch.is_tensor(opt.step) else torch.tensor(opt.step),
                   ema=opt.ema_grad_hand_pose.clone())
        if reset_mask is not None:
            out["reset_mask"] = reset_mask.clone()
    finally:
        mod.torch = real
    return out


def pk_chain_from_pack(pack, dtype=torch.float32):
    """Rebuilds the reference's pk.Chain from the frame tables of a hand pack (the tables were extracted from the URDF
    by tools/build_hand_pack.py; pk's own URDF loader needs lxml + transformations, absent here)."""
    ref = load()
    frame, tf = ref["pk_frame"], ref["pk_tf"]
    n = len(pack["frame_parent"])
    jt = {0: "fixed", 1: "revolute", 2: "prismatic"}

    def t3d(m34):
        m = torch.eye(4, dtype=dtype)
        m[:3] = torch.as_tensor(m34, dtype=dtype)
        return tf.Transform3d(matrix=m.unsqueeze(0), dtype=dtype)

    frames = []
    for f in range(n):
        j = frame.Joint(name=f"j{f}", offset=t3d(pack["frame_joff"][f]), joint_type=jt[int(pack["frame_jtype"][f])],
                        axis=[float(v) for v in pack["frame_axis"][f]], dtype=dtype)
        l = frame.Link(name=f"l{f}", offset=t3d(pack["frame_loff"][f]))
        frames.append(frame.Frame(name=f"f{f}", link=l, joint=j, children=[]))
    for f in range(n):
        p = int(pack["frame_parent"][f])
        if p >= 0:
            frames[p].children.append(frames[f])
    roots = [f for f in range(n) if int(pack["frame_parent"][f]) < 0]
    assert len(roots) == 1
    return ref["pk_chain"].Chain(frames[roots[0]], dtype=dtype), frames


# ---- the reference's OWN HandModel / ObjectModel methods on instances filled from a hand pack --------------------------
def load_models():
    """-> (HandModel, ObjectModel) classes of the reference.  Their modules import trimesh, pytorch3d and torchsdf at module
    level (absent here); those are only needed by the CONSTRUCTORS (URDF / mesh parsing, surface sampling), so they are
    stubbed and the constructors are never run: instances are made with __new__ and filled from the hand pack
    (`reference_hand` / `reference_object` below), then the reference's own methods run on them — cal_distance,
    self_penetration, get_surface_points, joint_entropy, pose_entropy, ObjectModel.cal_distance.  torchsdf.compute_sdf is
    oracle.hand.compute_sdf: the port of the TorchSDF kernel that tests/test_golden.py pins bit-exactly against the
    reference kernel body itself."""
    import importlib
    load()
    from oracle import hand as ohand
    for name in ("pytorch3d", "pytorch3d.ops", "pytorch3d.structures", "trimesh"):
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.modules["pytorch3d"].ops = sys.modules["pytorch3d.ops"]
    sys.modules["pytorch3d"].structures = sys.modules["pytorch3d.structures"]
    sdf = sys.modules.setdefault("torchsdf", types.ModuleType("torchsdf"))
    if not hasattr(sdf, "compute_sdf"):
        sdf.compute_sdf = ohand.compute_sdf
        sdf.index_vertices_by_faces = lambda v, f: v[f.long()]
    roma = sys.modules["roma"]
    if not hasattr(roma, "rotmat_to_rotvec"):  # pose_entropy only; restated (PARITY UNPINNED for that one call)
        from oracle import entropy as oent
        roma.rotmat_to_rotvec = oent.rotmat_to_rotvec
    os.environ.setdefault("SDF_BACKEND", "TORCHSDF")
    hm = importlib.import_module("graspqp.core.hand_model")
    om = importlib.import_module("graspqp.core.object_model")
    return hm.HandModel, om.ObjectModel


def reference_hand(pack, hand_pose, contact_idx=None, surface=None, dtype=torch.float32):
    """A reference HandModel instance (no constructor) in the state HandModel.set_parameters leaves it in
    (hand_model.py:833-873): hand_pose, global_translation / rotation, current_status = the reference chain's FK."""
    import collections
    HandModel, _ = load_models()
    ref = load()
    chain, frames = pk_chain_from_pack(pack, dtype)
    h = HandModel.__new__(HandModel)
    h.device = torch.device("cpu")
    t = lambda a: torch.as_tensor(np.asarray(a), dtype=dtype)
    names = [str(s) for s in pack["mesh_names"]]
    off = np.asarray(pack["face_offset"])
    fv = np.asarray(pack["face_verts"], np.float32).reshape(-1, 3, 3)
    h.mesh = collections.OrderedDict()
    for li, name in enumerate(names):
        m = {"face_verts": t(fv[off[li]:off[li + 1]]), "vertices": t(fv[off[li]:off[li + 1]].reshape(-1, 3)),
             "penetration_keypoints": t(pack["spheres"][pack["sphere_link"] == li][:, :3].reshape(-1, 3)),
             "contact_candidates": t(pack["cand"][pack["cand_link"] == li].reshape(-1, 3)),
             "normal_candidates": t(pack["cand_normal"][pack["cand_link"] == li].reshape(-1, 3))}
        if surface is not None:
            m["surface_points"] = t(surface[0][surface[1] == li].reshape(-1, 3))
        h.mesh[name] = m
    h.sphere_scales = t(pack["spheres"][np.argsort(pack["sphere_link"], kind="stable")][:, 3])
    h.penetration_keypoints_expanded = torch.zeros(1, len(pack["spheres"]), 3)   # read for its shape only (:1000-1007)
    h.global_index_to_link_index_penetration_expanded = None                     # read, never used
    h.joints_lower, h.joints_upper = t(pack["joints_lower"]), t(pack["joints_upper"])
    h.hand_pose = torch.as_tensor(hand_pose, dtype=dtype).clone().requires_grad_(True)
    h.global_translation = h.hand_pose[:, 0:3]
    h.global_rotation = ref["transforms"].robust_compute_rotation_matrix_from_ortho6d(h.hand_pose[:, 3:9])
    th = h.hand_pose[:, 9:]
    movable = [f for f in range(len(pack["frame_jtype"])) if int(pack["frame_jtype"][f]) != 0]
    gain = pack["frame_gain"] if "frame_gain" in pack else np.ones(len(pack["frame_parent"]), np.float32)
    assert "mlp_w1" not in pack, "use the hand's own joint_calc_fnc for the coupling network"
    res = chain.forward_kinematics({f"j{f}": th[:, int(pack["frame_dof"][f])] * float(gain[f]) for f in movable
                                    if int(pack["frame_dof"][f]) >= 0})
    h.current_status = {name: res[f"l{int(f)}"] for name, f in zip(names, pack["mesh_frame"])}
    return h


def reference_object(face_verts_list, batch_size_each, dtype=torch.float32):
    """A reference ObjectModel instance (no constructor) with what cal_distance reads (object_model.py:186-255)"""
    _, ObjectModel = load_models()
    o = ObjectModel.__new__(ObjectModel)
    o.device = torch.device("cpu")
    o.batch_size_each = batch_size_each
    o.sdf_library = "TORCHSDF"
    o.object_mesh_list = [None] * len(face_verts_list)
    o.object_face_verts_list = [torch.as_tensor(np.asarray(f, np.float32).reshape(-1, 3, 3)) for f in face_verts_list]
    o.object_scale_tensor = torch.ones(len(face_verts_list), batch_size_each, dtype=dtype)
    return o


def _hand_module(name):
    """graspqp/src/graspqp/hands/<name>.py loaded from its file (for its calculate_joints / calculate_jacobian); its
    `from graspqp.core import HandModel` is satisfied by the class load_models() imported"""
    import importlib.util
    HandModel, _ = load_models()
    core = sys.modules["graspqp.core"]
    had = hasattr(core, "HandModel")
    if not had:
        core.HandModel = HandModel
    try:
        spec = importlib.util.spec_from_file_location("ref_hand_" + name, os.path.join(REF, "graspqp", "src", "graspqp", "hands", name + ".py"))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
    finally:
        if not had:
            del core.HandModel
    return mod


def reference_hand_model(pack, dtype=torch.float32):
    """A reference HandModel instance made without its constructor, with everything HandModel.set_parameters, fk,
    _set_contact_idxs, get_contact_candidates, cal_distance, self_penetration and get_surface_points read
    (hand_model.py:762-1071), filled from the hand pack.  Links are keyed l<frame> (the names of pk_chain_from_pack's chain).
    The per-hand joint_calc_fnc is the reference's own (hands/panda.py, ability_hand.py, robotiq2.py)."""
    import collections
    HandModel, _ = load_models()
    chain, frames = pk_chain_from_pack(pack, dtype)
    h = HandModel.__new__(HandModel)
    h.device = torch.device("cpu")
    t = lambda a: torch.as_tensor(np.asarray(a), dtype=dtype)
    off = np.asarray(pack["face_offset"])
    fv = np.asarray(pack["face_verts"], np.float32).reshape(-1, 3, 3)
    h.mesh = collections.OrderedDict()
    for li, f in enumerate(pack["mesh_frame"]):
        h.mesh[f"l{int(f)}"] = {
            "face_verts": t(fv[off[li]:off[li + 1]]), "vertices": t(fv[off[li]:off[li + 1]].reshape(-1, 3)),
            "penetration_keypoints": t(pack["spheres"][pack["sphere_link"] == li][:, :3].reshape(-1, 3)),
            "contact_candidates": t(pack["cand"][pack["cand_link"] == li].reshape(-1, 3)),
            "normal_candidates": t(pack["cand_normal"][pack["cand_link"] == li].reshape(-1, 3))}
    order = np.argsort(pack["sphere_link"], kind="stable")
    h.sphere_scales = t(pack["spheres"][order][:, 3])
    h.penetration_keypoints = t(pack["spheres"][order][:, :3].reshape(-1, 3))
    h.global_index_to_link_index_penetration = torch.as_tensor(np.asarray(pack["sphere_link"])[order]).long()
    h.joints_lower, h.joints_upper = t(pack["joints_lower"]), t(pack["joints_upper"])
    h.n_dofs, h.n_contact_candidates = len(pack["joints_lower"]), len(pack["cand"])
    h.chain = chain
    name = str(pack["name"])
    h.joint_calc_fnc = _hand_module(name).calculate_joints if name in ("panda", "ability_hand", "robotiq2") else None
    movable = [f for f in range(len(pack["frame_jtype"])) if int(pack["frame_jtype"][f]) != 0]
    if "frame_joint" in pack:
        h.joints_names = [str(pack["frame_joint"][f]) for f in movable]  # robotiq2f140_fk reads it (utils/fk.py:41-47)
    # what get_contact_jacobian / get_req_joint_velocities read (hand_model.py:632-647, 772-777, 1079-1218)
    assert (np.diff(np.asarray(pack["cand_link"])) >= 0).all(), "candidates are stored in link order"
    h.global_index_to_link_index = torch.as_tensor(np.asarray(pack["cand_link"])).long()
    h.r_contact_body_frames = t(pack["cand"])
    mesh_frames = set(int(f) for f in pack["mesh_frame"])
    h._used_links_mask = [f in mesh_frames for f in range(len(pack["frame_parent"]))]  # chain frames in DFS order
    h.jacobian_fnc = _hand_module(name).calculate_jacobian if name in ("panda", "ability_hand", "robotiq2") else None
    h.hand_pose = h.contact_point_indices = h.contact_points = h.contact_normals = None
    h.dtype = dtype
    return h


def reference_object_model(face_verts_list, surface_points_list, batch_size_each, dtype=torch.float32):
    """reference_object + what calculate_energy reads of an ObjectModel (energy.py:16-62): cog, surface_points_tensor"""
    o = reference_object(face_verts_list, batch_size_each, dtype)
    sp = torch.stack([torch.as_tensor(s, dtype=dtype) for s in surface_points_list], 0)
    o.surface_points_tensor = sp.repeat_interleave(batch_size_each, dim=0)  # object_model.py:181-184
    o._cog = None                                                           # the reference's own property computes it (:64-68)
    o.dtype = dtype
    return o
