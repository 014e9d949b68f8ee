#!/usr/bin/env python
"""TEST INFRASTRUCTURE.  Generates tests/golden/*.npz by running the REFERENCE's own code in the build container:

  python tests/golden/make_golden.py          (needs /root/reference; see ref_harness.py for what is real and what is stubbed)

  sdf_ref.npz        thirdparty/TorchSDF forward kernel body, host-compiled from the reference source (oracle/_ref), on two
                     of TorchSDF's own test models (tests/models/cube.obj, sphere-42.obj) and seeded points
  fk_ref.npz         pytorch_kinematics Chain.forward_kinematics (+-0.01 rad perturbation included) and
                     Transform3d.transform_points / transform_normals on the Allegro tree
  metric_ref.npz     GraspSpanMetricFactory.create(GRASPQP) callable (span.py, registry.py, qp_solver.py) forward and
                     autograd backward; qpth replaced by oracle.qp (PARITY UNPINNED for the QP iterate)
  solver_kat.npz     graspqp/tests/metrics/test_solver.py's known-answer problem through SQPLsqSolver
  fit_ref_f32.npz    4 iterations of scripts/fit.py:381-458 with the reference MalaStar + calculate_energy, injected noise,
  fit_ref_f64.npz    run in fp32 (as the reference does) and in fp64 (the rounding-free target the kernels are held to)
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)
import ref_harness as rh  # noqa: E402
from graspqp_b200 import hand_pack, synth  # noqa: E402
from oracle import sdf as osdf  # noqa: E402


def load_obj(path):
    v, f = [], []
    for line in open(path):
        t = line.split()
        if not t:
            continue
        if t[0] == "v":
            v.append([float(x) for x in t[1:4]])
        elif t[0] == "f":
            f.append([int(x.split("/")[0]) - 1 for x in t[1:4]])
    v, f = np.asarray(v, np.float32), np.asarray(f, np.int64)
    return v[f]  # index_vertices_by_faces (torchsdf/sdf.py:5-31)


def gen_sdf():
    out = {}
    rng = np.random.default_rng(7)
    for name in ("cube", "sphere-42"):
        fv = load_obj(os.path.join(rh.REF, "thirdparty/TorchSDF/tests/models", name + ".obj"))
        pts = rng.uniform(-1, 1, (384, 3)).astype(np.float32)  # tests/value.py draws points in [-1,1]^3
        k = 96  # on / near vertices, edges and faces: the tie-break cases
        pts[:k] = fv.reshape(-1, 3)[rng.integers(0, len(fv) * 3, k)]
        pts[k:2 * k] = (fv[rng.integers(0, len(fv), k)].mean(1) * (1 + rng.normal(size=(k, 1)) * 0.01)).astype(np.float32)
        d2, sg, nr, cl = osdf.compute_sdf_ref(pts, fv)
        key = name.replace("-", "_")
        out.update({key + "_face_verts": fv, key + "_points": pts, key + "_dist2": d2, key + "_sign": sg,
                    key + "_normals": nr, key + "_clst": cl})
    np.savez_compressed(os.path.join(HERE, "sdf_ref.npz"), **out)
    print("sdf_ref.npz", {k: v.shape for k, v in out.items() if k.endswith("dist2")})


def gen_fk():
    out = {}
    for hand in ("allegro", "shadow_hand", "panda", "robotiq2"):
        pack = hand_pack.load_pack(hand)
        chain, frames = rh.pk_chain_from_pack(pack)
        rng = np.random.default_rng(3)
        B = 6
        lo, hi = pack["joints_lower"], pack["joints_upper"]
        theta = (lo + (hi - lo) * rng.random((B, len(lo)))).astype(np.float32)
        theta[0] = 0.0          # the perturbation's discontinuity: theta == 0 -> -0.01
        theta[1, ::2] = -0.05   # negative angles
        th = torch.tensor(theta)
        # hands/panda.py:6-14 feeds the single parameter to both finger joints: frame_dof repeats the column
        if hand != "robotiq2":
            th_dict = {f"j{f}": th[:, int(d)] for f, d in enumerate(pack["frame_dof"]) if d >= 0}
        else:
            # hands/robotiq2.py:9-13: the reference's own robotiq2f140_fk (clamp + its pickled network) expands the
            # parameter to the chain's joints, ordered like HandModel.joints_names (DFS over non-fixed joints)
            import importlib.util
            spec = importlib.util.spec_from_file_location(
                "ref_utils_fk", os.path.join(rh.REF, "graspqp", "src", "graspqp", "utils", "fk.py"))
            ref_fk = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(ref_fk)
            theta[2] = 0.9      # beyond the clamp (utils/fk.py:33)
            theta[3] = -0.2
            th = torch.tensor(theta)
            movable = [f for f in range(len(pack["frame_jtype"])) if int(pack["frame_jtype"][f]) != 0]
            with torch.no_grad():
                q = ref_fk.robotiq2f140_fk(th, joint_order=[str(pack["frame_joint"][f]) for f in movable])
            th_dict = {f"j{f}": q[:, k] for k, f in enumerate(movable)}
        res = chain.forward_kinematics(th_dict)
        link_m = np.stack([res[f"l{int(f)}"].get_matrix().expand(B, 4, 4).numpy() for f in pack["mesh_frame"]], 1)
        # contact candidates through Transform3d.transform_points / transform_normals (hand_model.py:1233-1248)
        cp, cn = [], []
        for li, f in enumerate(pack["mesh_frame"]):
            m = pack["cand_link"] == li
            if m.sum() == 0:
                continue
            T = res[f"l{int(f)}"]
            cp.append(T.transform_points(torch.tensor(pack["cand"][m])).expand(B, -1, -1))
            cn.append(T.transform_normals(torch.tensor(pack["cand_normal"][m])).expand(B, -1, -1))
        k = "" if hand == "allegro" else hand + "_"
        out.update({k + "theta": theta, k + "link_matrices": link_m.astype(np.float32),
                    k + "cand_points": torch.cat(cp, 1).numpy(), k + "cand_normals": torch.cat(cn, 1).numpy()})
        print("fk_ref", hand, link_m.shape)
    np.savez_compressed(os.path.join(HERE, "fk_ref.npz"), **out)


def gen_fk_more():
    """fk_more_ref.npz: pytorch_kinematics FK + contact transforms for the hands of SURVEY 8(f3): robotiq3 (plain chain) and
    ability_hand, whose 6 parameters are expanded to the chain's 10 joints by the reference's OWN get_all_joint_angles
    (hands/ability_hand.py:10-29, loaded from its file; its module-level `from graspqp.core import HandModel` is satisfied by
    a stub: HandModel itself needs trimesh / pytorch3d)."""
    import importlib.util
    import types
    out = {}
    for hand in ("robotiq3", "ability_hand"):
        pack = hand_pack.load_pack(hand)
        chain, frames = rh.pk_chain_from_pack(pack)
        rng = np.random.default_rng(4)
        B = 6
        lo, hi = pack["joints_lower"], pack["joints_upper"]
        theta = (lo + (hi - lo) * rng.random((B, len(lo)))).astype(np.float32)
        theta[0] = 0.0
        theta[1, ::2] = -0.05
        th = torch.tensor(theta)
        movable = [f for f in range(len(pack["frame_jtype"])) if int(pack["frame_jtype"][f]) != 0]
        if hand == "ability_hand":
            rh.load()
            core = sys.modules.get("graspqp.core")
            had = hasattr(core, "HandModel")
            if not had:
                core.HandModel = type("HandModel", (), {})
            spec = importlib.util.spec_from_file_location(
                "ref_ability_hand", os.path.join(rh.REF, "graspqp", "src", "graspqp", "hands", "ability_hand.py"))
            mod = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(mod)
            if not had:
                del core.HandModel
            q = mod.get_all_joint_angles(th)  # (B, 10) in chain order
            assert q.shape[1] == len(movable)
            th_dict = {f"j{f}": q[:, k] for k, f in enumerate(movable)}
        else:
            th_dict = {f"j{f}": th[:, int(pack["frame_dof"][f])] for f in movable}
        res = chain.forward_kinematics(th_dict)
        link_m = np.stack([res[f"l{int(f)}"].get_matrix().expand(B, 4, 4).numpy() for f in pack["mesh_frame"]], 1)
        cp, cn = [], []
        for li, f in enumerate(pack["mesh_frame"]):
            m = pack["cand_link"] == li
            if m.sum() == 0:
                continue
            T = res[f"l{int(f)}"]
            cp.append(T.transform_points(torch.tensor(pack["cand"][m])).expand(B, -1, -1))
            cn.append(T.transform_normals(torch.tensor(pack["cand_normal"][m])).expand(B, -1, -1))
        k = hand + "_"
        out.update({k + "theta": theta, k + "link_matrices": link_m.astype(np.float32),
                    k + "cand_points": torch.cat(cp, 1).numpy(), k + "cand_normals": torch.cat(cn, 1).numpy()})
        print("fk_more_ref", hand, link_m.shape)
    np.savez_compressed(os.path.join(HERE, "fk_more_ref.npz"), **out)


def gen_metric():
    ref = rh.load()
    out = {}
    for n_c in (12, 16, 4):
        rng = np.random.default_rng(100 + n_c)
        B = 5
        nrm = rng.normal(size=(B, n_c, 3)); nrm /= np.linalg.norm(nrm, axis=-1, keepdims=True)
        nrm[0, 0] = np.array([1.0, 1.0, 1.0]) / np.sqrt(3)  # the b1 switch of span.py:271-277
        cp = (-nrm * 0.08 + rng.normal(size=(B, n_c, 3)) * 0.01)
        cog = rng.normal(size=(B, 3)) * 0.005
        # the stored inputs are fp32: round first so that every run sees exactly the stored values
        nrm, cp, cog = [a.astype(np.float32).astype(np.float64) for a in (nrm, cp, cog)]
        for dt, tag in ((torch.float64, "f64"), (torch.float32, "f32")):
            F = ref["registry"].GraspSpanMetricFactory
            fn = F.create(F.MetricType.GRASPQP, solver_kwargs={"friction": 0.2, "max_limit": 20.0, "n_cone_vecs": 4})
            cpt = torch.tensor(cp, dtype=dt, requires_grad=True)
            nt = torch.tensor(nrm, dtype=dt)
            if dt == torch.float64:
                torch.set_default_dtype(torch.float64)  # the reference builds its eye()/ones() in the default dtype
            try:
                E, x = fn(contact_pts=cpt, contact_normals=nt, sdf=torch.zeros(B, n_c, dtype=dt),
                          cog=torch.tensor(cog, dtype=dt), with_solution=True, svd_gain=0.1)
                E.sum().backward()
            finally:
                torch.set_default_dtype(torch.float32)
            out.update({f"nc{n_c}_{tag}_E_fc": E.detach().numpy(), f"nc{n_c}_{tag}_x": x.detach().numpy(),
                        f"nc{n_c}_{tag}_grad_cpts": cpt.grad.numpy()})
        out.update({f"nc{n_c}_cpts": cp.astype(np.float32), f"nc{n_c}_nrm": nrm.astype(np.float32),
                    f"nc{n_c}_cog": cog.astype(np.float32)})
    # friction cone + grasp matrix alone (no solver behind them)
    m = ref["span"].OverallFrictionConeSpanMetric.from_dim(12, 6, batch_size=5, device="cpu", solver_cls=ref["qp_solver"].SQPLsqSolver,
                                                            friction=0.2, n_cone_vecs=4)
    cone = m.get_friction_cone(torch.tensor(out["nc12_nrm"]))
    out["nc12_cone"] = cone.numpy()
    np.savez_compressed(os.path.join(HERE, "metric_ref.npz"), **out)
    print("metric_ref.npz", sorted(out)[:6], "...")


def gen_solver_kat():
    ref = rh.load()
    A = torch.Tensor([[-0.0, 0.25819889, 0.25819889, -0.25819889, -0.25819889, -0.25819889],
                      [0.31622777, -0.18257419, -0.18257419, -0.18257419, -0.18257419, -0.18257419],
                      [0.02810913, -0.09128826, 0.10345754, 0.13997471, -0.05477109, -0.10345754]]).unsqueeze(0)
    b = torch.Tensor([0.25819889, 0.18257419, 0.00608464]).unsqueeze(0)
    S = ref["qp_solver"].SQPLsqSolver
    solver = S.from_mat(A, b)
    value, x = solver(A, b, min_bound=-10.0, max_bound=1e3, init=0.1, return_solution=True)
    assert torch.allclose(value, torch.zeros_like(value), atol=1e-4), value  # graspqp/tests/metrics/test_solver.py:26
    np.savez_compressed(os.path.join(HERE, "solver_kat.npz"), A=A.numpy(), b=b.numpy(), value=value.numpy(), x=x.numpy())
    print("solver_kat.npz value", value.numpy())


def gen_fit(dtype=torch.float32, name="fit_ref_f32.npz"):
    pack = hand_pack.load_pack("allegro")
    B, n_c, n_pts, T, seed = 6, 12, 192, 4, 3
    v, f = synth.make_mesh("sphere", 400, seed)
    fv = synth.face_vertices(v, f)
    sp = synth.sample_surface(v, f, n_pts, seed=seed)
    pose, idx = synth.random_grasp_init(pack, [v], B, n_c, seed)
    pose[:, :3] *= 0.7
    if dtype == torch.float64:
        torch.set_default_dtype(torch.float64)  # the reference builds eye()/ones() in the default dtype
    st = rh.make_fit_state(pack, [fv], [sp], B, pose.astype(np.float32), idx, dtype=dtype)
    out = dict(face_verts=fv, surface=sp.astype(np.float32), pose0=pose.astype(np.float32), idx0=idx.astype(np.int32),
               energy0=st.energy.detach().numpy().copy(), first_grad=st.first_grad.numpy().copy())
    for k, v_ in st.losses.items():
        out["loss0_" + k] = v_.detach().numpy().copy()  # st.losses is updated in place later
    rng = np.random.default_rng(0)
    n_cand = len(pack["cand"])
    for t in range(T):
        us, ir, ua = rng.random((B, n_c), dtype=np.float32), rng.integers(0, n_cand, (B, n_c)).astype(np.int32), rng.random(B, dtype=np.float32)
        o = rh.fit_iterate(st, torch.tensor(us), torch.tensor(ir).long(), torch.tensor(ua))
        out.update({f"t{t}_u_switch": us, f"t{t}_idx_rand": ir, f"t{t}_u_accept": ua})
        for k in ("s", "z", "proposal", "idx_prop", "new_energy", "new_grad", "accept", "temperature", "energy", "pose", "idx"):
            out[f"t{t}_{k}"] = o[k].numpy()
        for k, v_ in o["new_losses"].items():
            out[f"t{t}_loss_{k}"] = v_.numpy()
    torch.set_default_dtype(torch.float32)
    np.savez_compressed(os.path.join(HERE, name), **out)
    print(name, "accepts", [out[f"t{t}_accept"].sum() for t in range(T)])


def gen_fit_reset(name="fit_reset_ref_f64.npz"):
    """M3: iterations around two RESET iterations (fit.py:408-422) with the reference's own MalaStar (clip_grad=True:
    optimizer.py:211-213), a NaN gradient entry injected before iteration 1, the reset mask once derived from the z-score
    (threshold 0.3 so that some rows are masked and some are not) and once injected; initialize_convex_hull's output
    is injected (a second random grasp initialisation)."""
    dtype = torch.float64
    pack = hand_pack.load_pack("allegro")
    B, n_c, n_pts, seed = 6, 12, 192, 3
    v, f = synth.make_mesh("sphere", 400, seed)
    fv = synth.face_vertices(v, f)
    sp = synth.sample_surface(v, f, n_pts, seed=seed)
    pose, idx = synth.random_grasp_init(pack, [v], B, n_c, seed)
    pose[:, :3] *= 0.7
    torch.set_default_dtype(torch.float64)
    st = rh.make_fit_state(pack, [fv], [sp], B, pose.astype(np.float32), idx, dtype=dtype, clip_grad=True)
    out = dict(face_verts=fv, surface=sp.astype(np.float32), pose0=pose.astype(np.float32), idx0=idx.astype(np.int32))
    rng = np.random.default_rng(7)
    n_cand = len(pack["cand"])
    plan = ["plain", "nan", "reset_z", "plain", "reset_mask", "plain"]
    out["plan"] = np.array(plan)
    for t, kind in enumerate(plan):
        us, ir, ua = rng.random((B, n_c), dtype=np.float32), rng.integers(0, n_cand, (B, n_c)).astype(np.int32), rng.random(B, dtype=np.float32)
        reset = None
        if kind == "nan":
            st.hand.hand_pose.grad[1, 4] = float("nan")
            st.hand.hand_pose.grad[2, 11] = 1e4  # beyond the clip
        if kind.startswith("reset"):
            pi, ii = synth.random_grasp_init(pack, [v], B, n_c, 50 + t)
            pi[:, :3] *= 0.8
            mask = None if kind == "reset_z" else torch.tensor([True, False, False, True, False, False])
            reset = (torch.tensor(pi.astype(np.float32)), torch.tensor(ii), mask)
            out.update({f"t{t}_pose_init": pi.astype(np.float32), f"t{t}_idx_init": ii.astype(np.int32)})
            if mask is not None:
                out[f"t{t}_mask_in"] = mask.numpy()
        o = rh.fit_iterate(st, torch.tensor(us), torch.tensor(ir).long(), torch.tensor(ua), reset=reset,
                           z_score_threshold=0.3)
        out.update({f"t{t}_u_switch": us, f"t{t}_idx_rand": ir, f"t{t}_u_accept": ua})
        for k in ("s", "z", "proposal", "idx_prop", "new_energy", "new_grad", "accept", "temperature", "energy", "pose",
                  "idx", "step", "ema") + (("reset_mask",) if reset is not None else ()):
            out[f"t{t}_{k}"] = o[k].numpy()
        for k, v_ in o["new_losses"].items():
            out[f"t{t}_loss_{k}"] = v_.numpy()
        if reset is not None:
            print("  reset iteration", t, "mask", o["reset_mask"].numpy().astype(int))
    torch.set_default_dtype(torch.float32)
    np.savez_compressed(os.path.join(HERE, name), **out)
    print(name, "accepts", [int(out[f"t{t}_accept"].sum()) for t in range(len(plan))])


def gen_fit_alt(name="fit_alt_ref.npz"):
    """(f3) scripts/fit.py:381-458 with the reference's OTHER optimiser and energies, injected noise; "dexanneal" in fp32
    (calc_e_fc is a TorchScript function with float constants, dexgrasp.py:18-30: it rejects doubles), "tdgmala" in fp64:
       run "dexanneal": --optimizer dexgraspnet (AnnealingDexGraspNet) + --energy_type dexgrasp, 5 iterations, the 3rd a
                        reset iteration with an injected mask (resets are neither force-accepted nor seen by this optimiser);
                        stepsize_period 2 / annealing_period 3 so that both schedules move within the run;
       run "tdgmala":   MalaStar + --energy_type tdg (1000 directions from numpy seed 7), 3 iterations;
       run "priorwall": MalaStar + graspqp with the optional terms E_prior / E_wall on;
       run "malareset": MalaStar + graspqp, a reset iteration followed by one with forced rejections (the gradient that
                        rejected rows keep there is old + new: oracle/pipeline.py AFTER_RESET)."""
    pack = hand_pack.load_pack("allegro")
    B, n_c, n_pts, seed = 6, 12, 192, 3
    v, f = synth.make_mesh("sphere", 400, seed)
    fv = synth.face_vertices(v, f)
    sp = synth.sample_surface(v, f, n_pts, seed=seed)
    pose, idx = synth.random_grasp_init(pack, [v], B, n_c, seed)
    pose[:, :3] *= 0.7
    out = dict(face_verts=fv, surface=sp.astype(np.float32), pose0=pose.astype(np.float32), idx0=idx.astype(np.int32))
    n_cand = len(pack["cand"])
    try:
        for run, dtype, kw, plan in (("dexanneal", torch.float32,
                                      dict(energy_type="dexgrasp", optimizer="dexgraspnet", switch_possibility=0.5,
                                           stepsize_period=2, annealing_period=3),
                                      ["plain", "plain", "reset_mask", "plain", "plain"]),
                                     ("tdgmala", torch.float64, dict(energy_type="tdg"), ["plain", "plain", "plain"]),
                                     # MalaStar + graspqp around a reset iteration, with rejections forced in the iteration
                                     # AFTER it (u_accept 1e30 on two rows): oracle.pipeline.AFTER_RESET
                                     ("malareset", torch.float64, dict(), ["plain", "reset_mask", "reject", "plain"]),
                                     # the reference's energy.py:68-77 with --w_prior 5 --w_wall 50 (HandOracle supplies
                                     # global_rotation, grasp_axis and get_surface_points over the stored sample table)
                                     ("priorwall", torch.float64, dict(weights=dict(E_prior=5.0, E_wall=50.0),
                                                                       hand_surface=hand_pack.surface_points(pack, 128)),
                                      ["plain", "plain"])):
            torch.set_default_dtype(dtype)
            st = rh.make_fit_state(pack, [fv], [sp], B, pose.astype(np.float32), idx, dtype=dtype, **kw)
            if run == "priorwall":
                out["hand_surf"], out["hand_surf_link"] = kw["hand_surface"]
            if run == "tdgmala":
                out["tdg_dirs"] = st.energy_fnc.tdg_energy.target_direction_6D[0, :, :3].numpy().astype(np.float32)
            rng = np.random.default_rng(11)
            out[run + "_plan"] = np.array(plan)
            out[run + "_energy0"] = st.energy.detach().numpy().copy()
            out[run + "_first_grad"] = st.first_grad.numpy().copy()
            for t, kind in enumerate(plan):
                us, ir, ua = rng.random((B, n_c), dtype=np.float32), rng.integers(0, n_cand, (B, n_c)).astype(np.int32), rng.random(B, dtype=np.float32)
                reset = None
                k0 = f"{run}_t{t}_"
                if kind == "reject":
                    ua[[1, 4]] = 1e30  # outside [0,1): rows 1 and 4 are rejected whatever their energies
                if kind == "reset_mask":
                    pi, ii = synth.random_grasp_init(pack, [v], B, n_c, 70 + t)
                    pi[:, :3] *= 0.8
                    mask = torch.tensor([True, False, True, True, False, False])
                    reset = (torch.tensor(pi.astype(np.float32)), torch.tensor(ii), mask)
                    out.update({k0 + "pose_init": pi.astype(np.float32), k0 + "idx_init": ii.astype(np.int32),
                                k0 + "mask_in": mask.numpy()})
                o = rh.fit_iterate(st, torch.tensor(us), torch.tensor(ir).long(), torch.tensor(ua), reset=reset)
                out.update({k0 + "u_switch": us, k0 + "idx_rand": ir, k0 + "u_accept": ua})
                for k in ("s", "z", "proposal", "idx_prop", "new_energy", "new_grad", "accept", "temperature", "energy",
                          "pose", "idx", "step", "ema"):
                    out[k0 + k] = o[k].numpy()
                for k, v_ in o["new_losses"].items():
                    out[k0 + "loss_" + k] = v_.numpy()
            print(name, run, "accepts", [int(out[f"{run}_t{t}_accept"].sum()) for t in range(len(plan))],
                  "E_fc", out[f"{run}_t0_loss_E_fc"][:3])
    finally:
        torch.set_default_dtype(torch.float32)
    np.savez_compressed(os.path.join(HERE, name), **out)


MODEL_HANDS = ("allegro", "shadow_hand", "panda", "ability_hand", "robotiq3")


def models_case(hand, seed=3):
    """inputs of gen_models / the tests that read models_ref.npz: poses close to a 400-face sphere with the joints pushed towards
    their upper limits for half of the rows (self-penetration active), object surface samples, hand surface samples"""
    pack = hand_pack.load_pack(hand)
    v, f = synth.make_mesh("sphere", 400, seed)
    B = 4
    pose, idx = synth.random_grasp_init(pack, [v], B, 4, seed)
    pose[:, :3] *= 0.6
    lo, hi = pack["joints_lower"], pack["joints_upper"]
    pose[B // 2:, 9:] = (lo + 0.92 * (hi - lo))[None]
    sp = synth.sample_surface(v, f, 96, seed=seed)
    return pack, synth.face_vertices(v, f), pose.astype(np.float32), idx, sp.astype(np.float32), hand_pack.surface_points(pack, 64)


def gen_models():
    """models_ref.npz: the reference's OWN HandModel.cal_distance / self_penetration / get_surface_points / joint_entropy /
    pose_entropy and ObjectModel.cal_distance (forward + autograd backward), run on instances made without their constructors
    and filled from the hand packs (ref_harness.reference_hand / reference_object).  fp32, as the reference runs."""
    HandModel, ObjectModel = rh.load_models()
    out = {}
    for hand in MODEL_HANDS:
        pack, fv, pose, idx, sp, surf = models_case(hand)
        B = len(pose)
        h = rh.reference_hand(pack, pose, surface=surf)
        x = torch.tensor(sp).unsqueeze(0).repeat(B, 1, 1)
        d = HandModel.cal_distance(h, x)
        e = HandModel.self_penetration(h)
        s = HandModel.get_surface_points(h)
        (torch.relu(d).sum() + e.sum() + s[..., 2].clamp(max=0).abs().sum()).backward()
        te, re_ = HandModel.pose_entropy(h)
        k = hand + "_"
        out.update({k + "hand_distance": d.detach().numpy(), k + "e_spen": e.detach().numpy(),
                    k + "surface_points": s.detach().numpy(), k + "grad": h.hand_pose.grad.numpy().copy(),
                    k + "entropy": np.array([float(HandModel.joint_entropy(h)), float(te), float(re_)], np.float32)})
        # ObjectModel.cal_distance at this hand's contact candidates' world positions (any points would do)
        o = rh.reference_object([fv], B)
        pts = s[:, :12].detach().clone().requires_grad_(True)
        dist, nrm = ObjectModel.cal_distance(o, pts)
        dist.abs().sum().backward()
        out.update({k + "obj_points": pts.detach().numpy(), k + "obj_distance": dist.detach().numpy(),
                    k + "obj_normals": nrm.detach().numpy(), k + "obj_grad": pts.grad.numpy().copy()})
        print("models_ref", hand, "max depth", float(d.max()), "E_spen", e.detach().numpy().round(4))
    np.savez_compressed(os.path.join(HERE, "models_ref.npz"), **out)


def gen_metric_alt():
    """metric_alt_ref.npz: the reference's DexgraspSpanMetric (metrics/ops/dexgrasp.py) and TDGSpanMetric (metrics/ops/tdg.py)
    forward + autograd backward in fp32 (TDGEnergy holds fp32 constants), with the 1000 target directions TDGEnergy drew."""
    ref = rh.load()
    reg = ref["registry"]
    out = {}
    for n_c in (12, 4):
        rng = np.random.default_rng(200 + n_c)
        B = 6
        nrm = rng.normal(size=(B, n_c, 3)); nrm /= np.linalg.norm(nrm, axis=-1, keepdims=True)
        nrm[0, 0] = np.array([0.05, 0.998, 0.02]) / np.linalg.norm([0.05, 0.998, 0.02])  # the base switch of tdg.py:101-102
        nrm[1] = nrm[1, :1] + rng.normal(size=(n_c, 3)) * 0.05                            # one-sided grasp: density != 1/1
        nrm[1] /= np.linalg.norm(nrm[1], axis=-1, keepdims=True)
        cp = (-nrm * 0.08 + rng.normal(size=(B, n_c, 3)) * 0.01)
        cog = rng.normal(size=(B, 3)) * 0.005
        nrm, cp, cog = [a.astype(np.float32) for a in (nrm, cp, cog)]
        k = f"nc{n_c}_"
        out.update({k + "cpts": cp, k + "nrm": nrm, k + "cog": cog})
        for tw in (0.0, 5.0):
            cpt = torch.tensor(cp, requires_grad=True)
            E, _ = reg.DexgraspSpanMetric()(cpt, torch.tensor(nrm), torch.tensor(cog), torque_weight=tw, with_solution=True)
            E.sum().backward()
            out.update({k + f"dex_tw{tw:g}_E": E.detach().numpy(), k + f"dex_tw{tw:g}_grad": cpt.grad.numpy()})
        np.random.seed(7)  # scripts/fit.py:159 seeds numpy before the metric is built
        m = reg.TDGSpanMetric()
        out[k + "tdg_dirs"] = m.tdg_energy.target_direction_6D[0, :, :3].numpy().copy()
        cpt = torch.tensor(cp, requires_grad=True)
        E, _ = m(cpt, torch.tensor(nrm), torch.tensor(cog))
        E.sum().backward()
        out.update({k + "tdg_E": E.detach().numpy(), k + "tdg_grad": cpt.grad.numpy()})
        print("metric_alt", n_c, "dex", out[k + "dex_tw5_E"][:3], "tdg", out[k + "tdg_E"][:3])
    np.savez_compressed(os.path.join(HERE, "metric_alt_ref.npz"), **out)


if __name__ == "__main__":
    assert rh.available(), "reference not present"
    import subprocess
    subprocess.check_call(["bash", os.path.join(ROOT, "oracle", "build_ref.sh")], stdout=subprocess.DEVNULL)
    gen_sdf(); gen_fk(); gen_solver_kat(); gen_metric(); gen_fit(); gen_fit(torch.float64, "fit_ref_f64.npz"); gen_fit_reset(); gen_metric_alt(); gen_fit_alt(); gen_fk_more(); gen_models()
