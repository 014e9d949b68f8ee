"""CPU suite against the committed golden vectors (outputs of the reference's own code, tests/golden/make_golden.py):
the oracle restatements, and the kernel sources run by the host emulator."""
import os

import numpy as np
import pytest
import torch

import golden_cases as gc
from graspqp_b200 import engine, hand_pack
from oracle import qp as oqp
from oracle import sdf as osdf
from oracle.hand import HandOracle
from oracle.pipeline import FitLoopOracle


def test_oracle_sdf_port_matches_reference_kernel(oracle_built):
    g = gc.load("sdf_ref")
    for key in ("cube", "sphere_42"):
        got = osdf.compute_sdf_port(g[key + "_points"], g[key + "_face_verts"])
        for a, name in zip(got, ("_dist2", "_sign", "_normals", "_clst")):
            np.testing.assert_array_equal(a, g[key + name], err_msg=key + name)
        # thirdparty/TorchSDF/tests/normal.py:36-39: normal * 2 sqrt(d2) == d(d2)/dp = 2 (p - c)
        nr, d2, cl = g[key + "_normals"], g[key + "_dist2"], g[key + "_clst"]
        np.testing.assert_allclose(nr * 2 * np.sqrt(d2)[:, None], 2 * (g[key + "_points"] - cl), atol=5e-7)


@pytest.mark.parametrize("hand", ["allegro", "shadow_hand", "panda", "robotiq2", "robotiq3", "ability_hand"])
def test_oracle_fk_matches_pytorch_kinematics(hand):
    g = gc.load("fk_more_ref" if hand in gc.MORE_HANDS else "fk_ref")
    k = "" if hand == "allegro" else hand + "_"
    pack = hand_pack.load_pack(hand)
    o = HandOracle(pack, torch.float32)
    st = o.fk(torch.tensor(g[k + "theta"]))
    got = np.stack([st[int(f)].numpy() for f in pack["mesh_frame"]], 1)
    np.testing.assert_allclose(got, g[k + "link_matrices"], atol=1e-6)


def test_oracle_solver_known_answer():
    """graspqp/tests/metrics/test_solver.py:5-26 (residual ~ 0, atol 1e-4) and the converged optimum from scipy
    (the reference's ScipyLsqSolver, metrics/solver/scipy_solver.py:64-125)."""
    from scipy.optimize import lsq_linear
    g = gc.load("solver_kat")
    A, b = torch.tensor(g["A"]), torch.tensor(g["b"])
    value, x = oqp.lsq_solve(A, b, -10.0, 1e3)
    assert torch.allclose(value, torch.zeros_like(value), atol=1e-4)
    np.testing.assert_allclose(value.numpy(), g["value"], atol=1e-6)
    r = lsq_linear(g["A"][0].astype(np.float64), g["b"][0].astype(np.float64), bounds=(-10.0, 1e3))
    assert 0.5 * np.sum((g["A"][0] @ r.x - g["b"][0]) ** 2) < 1e-4


def test_oracle_metric_matches_reference_metric():
    g = gc.load("metric_ref")
    cone = oqp.friction_cone(torch.tensor(g["nc12_nrm"]))
    np.testing.assert_allclose(cone.numpy(), g["nc12_cone"], rtol=1e-6, atol=1e-7)
    for n_c in (12, 16, 4):
        t = lambda k: torch.tensor(g[f"nc{n_c}_{k}"], dtype=torch.float64)
        cp = t("cpts").requires_grad_(True)
        E, lam, _ = oqp.graspqp_energy(cp, t("nrm"), t("cog"))
        E.sum().backward()
        np.testing.assert_allclose(E.detach().numpy(), g[f"nc{n_c}_f64_E_fc"], rtol=1e-9)
        np.testing.assert_allclose(lam.detach().numpy(), g[f"nc{n_c}_f64_x"], rtol=1e-8, atol=1e-10)
        np.testing.assert_allclose(cp.grad.numpy(), g[f"nc{n_c}_f64_grad_cpts"], rtol=1e-7, atol=1e-9)


def test_oracle_loop_matches_reference_loop():
    """oracle.pipeline (our restatement of optimizer.py / energy.py / fit.py glue) vs the reference's own MalaStar +
    calculate_energy on the same injected noise, both fp32 on the CPU."""
    g = gc.load("fit_ref_f32")
    pack = hand_pack.load_pack("allegro")
    B = len(g["pose0"])
    orc = FitLoopOracle(pack, [g["face_verts"]], [g["surface"]], B, g["pose0"], g["idx0"].astype(np.int64), dtype=torch.float32)
    np.testing.assert_allclose(orc.energy.numpy(), g["energy0"], rtol=1e-5)
    np.testing.assert_allclose(orc.first_grad.numpy(), g["first_grad"], rtol=1e-3, atol=1e-3 * np.abs(g["first_grad"]).max())
    for t in range(4):
        out = orc.iterate(torch.tensor(g[f"t{t}_u_switch"]), torch.tensor(g[f"t{t}_idx_rand"]).long(),
                          torch.tensor(g[f"t{t}_u_accept"]))
        np.testing.assert_array_equal(out["idx_prop"].numpy(), g[f"t{t}_idx_prop"])
        np.testing.assert_allclose(out["proposal"].numpy(), g[f"t{t}_proposal"], rtol=1e-5, atol=1e-6)
        np.testing.assert_allclose(out["new_energy"].numpy(), g[f"t{t}_new_energy"], rtol=1e-5)
        np.testing.assert_array_equal(out["accept"].numpy(), g[f"t{t}_accept"])
        np.testing.assert_allclose(out["temperature"].numpy(), g[f"t{t}_temperature"], rtol=1e-6)
        np.testing.assert_allclose(orc.energy.numpy(), g[f"t{t}_energy"], rtol=1e-5)


def test_oracle_reset_iterations_match_reference():
    """oracle.pipeline's reset path (MalaStarOracle.reset_envs, HandOracle.set_parameters(env_mask)) and clip_grad / NaN
    handling vs the reference's own MalaStar.reset_envs / try_step (fit_reset_ref_f64.npz), fp64 on both sides."""
    g = gc.load("fit_reset_ref_f64")
    pack = hand_pack.load_pack("allegro")
    B = len(g["pose0"])
    orc = FitLoopOracle(pack, [g["face_verts"]], [g["surface"]], B, g["pose0"], g["idx0"].astype(np.int64),
                        dtype=torch.float64, clip_grad=True)
    t64 = lambda a: torch.tensor(a, dtype=torch.float64)
    for t, kind in enumerate(g["plan"]):
        if kind == "nan":
            orc.hand.hand_pose.grad[1, 4] = float("nan")
            orc.hand.hand_pose.grad[2, 11] = 1e4
        reset = None
        if str(kind).startswith("reset"):
            mask = g[f"t{t}_mask_in"] if f"t{t}_mask_in" in g else None
            reset = (g[f"t{t}_pose_init"], g[f"t{t}_idx_init"], mask)
        out = orc.iterate(t64(g[f"t{t}_u_switch"]), torch.tensor(g[f"t{t}_idx_rand"]).long(), t64(g[f"t{t}_u_accept"]),
                          reset=reset, z_score_threshold=0.3)
        if reset is not None:
            np.testing.assert_array_equal(out["reset_mask"].numpy(), g[f"t{t}_reset_mask"])
        np.testing.assert_array_equal(out["idx_prop"].numpy(), g[f"t{t}_idx_prop"])
        # the reference keeps mu / step_size / temperature_decay as float32 tensors (optimizer.py:176-181) even in an
        # fp64 run, the oracle as doubles: 5e-8 relative in the proposal, amplified by the energy's curvature
        np.testing.assert_allclose(out["proposal"].numpy(), g[f"t{t}_proposal"], rtol=1e-6, atol=1e-9)
        np.testing.assert_allclose(out["new_energy"].numpy(), g[f"t{t}_new_energy"], rtol=1e-5)
        gw = g[f"t{t}_new_grad"]
        np.testing.assert_allclose(out["new_grad"].numpy(), gw, rtol=1e-4, atol=1e-5 * np.abs(gw).max())
        np.testing.assert_array_equal(out["accept"].numpy(), g[f"t{t}_accept"])
        np.testing.assert_allclose(orc.energy.numpy(), g[f"t{t}_energy"], rtol=1e-5)
        np.testing.assert_array_equal(orc.opt.step.numpy(), g[f"t{t}_step"])
        np.testing.assert_allclose(orc.opt.ema.numpy(), g[f"t{t}_ema"], rtol=1e-5, atol=1e-12)


def test_oracle_other_energies_match_reference():
    """oracle.metrics_alt (dexgrasp, tdg) vs the reference's DexgraspSpanMetric / TDGSpanMetric (metric_alt_ref.npz, fp32 runs)"""
    from oracle import metrics_alt as ma
    g = gc.load("metric_alt_ref")
    for n_c in (12, 4):
        k = f"nc{n_c}_"
        for dt in (torch.float32, torch.float64):
            cp = torch.tensor(g[k + "cpts"], dtype=dt, requires_grad=True)
            n, cog = torch.tensor(g[k + "nrm"], dtype=dt), torch.tensor(g[k + "cog"], dtype=dt)
            tdg = ma.TDGOracle(g[k + "tdg_dirs"])
            for key, fn in (("dex_tw0", lambda: ma.dexgrasp_energy(cp, n, cog, 0.0)),
                            ("dex_tw5", lambda: ma.dexgrasp_energy(cp, n, cog, 5.0)), ("tdg", lambda: tdg.energy(cp, n, cog))):
                cp.grad = None
                E = fn()
                E.sum().backward()
                np.testing.assert_allclose(E.detach().numpy(), g[k + key + "_E"], rtol=2e-6, err_msg=key)
                want = g[k + key + "_grad"]
                got = cp.grad.numpy() if cp.grad is not None else np.zeros_like(want)
                np.testing.assert_allclose(got, want, rtol=0, atol=2e-6 * max(np.abs(want).max(), 1e-30), err_msg=key)


def test_oracle_other_optimizer_and_energies_match_reference_loop():
    """oracle.pipeline with AnnealingOracle / dexgrasp / tdg and the after-reset gradient rule vs the reference's own
    AnnealingDexGraspNet, metrics and MalaStar in scripts/fit.py's loop body (fit_alt_ref.npz)."""
    g = gc.load("fit_alt_ref")
    pack = hand_pack.load_pack("allegro")
    B = len(g["pose0"])
    for run, dt in (("dexanneal", torch.float32), ("tdgmala", torch.float64), ("malareset", torch.float64),
                    ("priorwall", torch.float64)):
        kw = dict(gc.ALT_RUNS[run])
        if run == "tdgmala":
            kw["tdg_directions"] = g["tdg_dirs"]
        if run == "priorwall":
            kw["hand_surface"] = (g["hand_surf"], g["hand_surf_link"])
        orc = FitLoopOracle(pack, [g["face_verts"]], [g["surface"]], B, g["pose0"], g["idx0"].astype(np.int64), dtype=dt, **kw)
        np.testing.assert_allclose(orc.energy.numpy(), g[run + "_energy0"], rtol=1e-6)
        T = lambda a: torch.tensor(a, dtype=dt)
        for t, kind in enumerate(g[run + "_plan"]):
            k = f"{run}_t{t}_"
            reset = (g[k + "pose_init"], g[k + "idx_init"], g[k + "mask_in"]) if kind == "reset_mask" else None
            out = orc.iterate(T(g[k + "u_switch"]), torch.tensor(g[k + "idx_rand"]).long(), T(g[k + "u_accept"]), reset=reset)
            np.testing.assert_array_equal(out["idx_prop"].numpy(), g[k + "idx_prop"])
            # the proposal of the iteration after a rejected row depends on the gradient that row kept: 1e-6 here means the
            # after-reset rule (old + new gradient) is the reference's
            np.testing.assert_allclose(out["proposal"].numpy(), g[k + "proposal"], rtol=1e-6, atol=1e-8, err_msg=k)
            np.testing.assert_allclose(out["new_energy"].numpy(), g[k + "new_energy"], rtol=1e-6, err_msg=k)
            gw = g[k + "new_grad"]
            np.testing.assert_allclose(out["new_grad"].numpy(), gw, rtol=1e-4, atol=1e-5 * np.abs(gw).max(), err_msg=k)
            np.testing.assert_array_equal(out["accept"].numpy(), g[k + "accept"])
            np.testing.assert_allclose(np.asarray(out["temperature"]), g[k + "temperature"], rtol=1e-6)
            np.testing.assert_allclose(orc.opt.ema.numpy(), g[k + "ema"], rtol=1e-5, atol=1e-12)


def test_emulated_kernels_other_energies(emu):
    gc.check_metric_alt(emu, engine.NumpyBackend())


def test_emulated_kernels_other_optimizer_and_energies(emu):
    """AnnealingDexGraspNet + dexgrasp, MalaStar + tdg, and the after-reset gradient rule through the emulated kernels"""
    assert gc.check_fit_alt(emu, engine.NumpyBackend()) == 3


@pytest.mark.parametrize("hand", gc.MODEL_HANDS)
def test_oracle_models_match_reference_methods(hand):
    """oracle.hand.HandOracle.cal_distance / self_penetration / get_surface_points, oracle.pipeline.ObjectOracle.cal_distance
    and oracle.entropy vs the reference's OWN HandModel / ObjectModel methods (models_ref.npz): forward bit-exact in fp32,
    autograd gradients to 1e-6 of the largest entry"""
    from oracle import entropy as oe
    from oracle.pipeline import ObjectOracle
    g = gc.load("models_ref")
    pack, fv, pose, idx, sp, surf = gc.models_case(hand)
    k = hand + "_"
    B = len(pose)
    o = HandOracle(pack, torch.float32)
    o.set_surface_points(*surf)
    o.set_parameters(torch.tensor(pose).requires_grad_(True), torch.tensor(idx).long())
    x = torch.tensor(sp).unsqueeze(0).repeat(B, 1, 1)
    d, e, s = o.cal_distance(x), o.self_penetration(), o.get_surface_points()
    np.testing.assert_array_equal(d.detach().numpy(), g[k + "hand_distance"])
    np.testing.assert_allclose(e.detach().numpy(), g[k + "e_spen"], rtol=1e-6, atol=1e-9)
    np.testing.assert_allclose(s.detach().numpy(), g[k + "surface_points"], atol=2e-8)
    (torch.relu(d).sum() + e.sum() + s[..., 2].clamp(max=0).abs().sum()).backward()
    gw = g[k + "grad"]
    np.testing.assert_allclose(o.hand_pose.grad.numpy(), gw, rtol=0, atol=1e-6 * np.abs(gw).max())
    je = float(oe.joint_entropy(o.hand_pose.detach(), o.joints_lower, o.joints_upper))
    te, re_ = [float(t) for t in oe.pose_entropy(o.hand_pose.detach())]
    np.testing.assert_allclose([je, te, re_], g[k + "entropy"], rtol=1e-6)
    obj = ObjectOracle([fv], [sp], B, torch.float32)
    pts = torch.tensor(g[k + "obj_points"]).requires_grad_(True)
    dist, nrm = obj.cal_distance(pts)
    dist.abs().sum().backward()
    np.testing.assert_array_equal(dist.detach().numpy(), g[k + "obj_distance"])
    np.testing.assert_array_equal(nrm.detach().numpy(), g[k + "obj_normals"])
    np.testing.assert_allclose(pts.grad.numpy(), g[k + "obj_grad"], rtol=1e-6, atol=1e-9)


def test_emulated_kernels_models(emu):
    """H4 / H5 / O1 kernels vs the reference's own HandModel.cal_distance / self_penetration / ObjectModel.cal_distance"""
    gc.check_models(emu, engine.NumpyBackend())


def test_emulated_kernels_sdf(emu):
    gc.check_sdf(emu, engine.NumpyBackend())


def test_emulated_kernels_fk(emu):
    gc.check_fk(emu, engine.NumpyBackend())


def test_emulated_kernels_fk_more_hands(emu):
    """(f3) robotiq3 and the ability hand (mimic joints: GqpHandDesc::frame_gain_host)"""
    gc.check_fk(emu, engine.NumpyBackend(), hands=gc.MORE_HANDS)


def test_emulated_kernels_metric(emu):
    gc.check_metric(emu, engine.NumpyBackend())


def test_emulated_kernels_fit_iteration(emu):
    assert gc.check_fit(emu, engine.NumpyBackend(), T=2) > 0


def test_emulated_kernels_reset_iterations(emu):
    """M3 + clip_grad / NaN through the emulated kernels against the reference's own MalaStar"""
    assert gc.check_fit_reset(emu, engine.NumpyBackend()) == 4


@pytest.mark.skipif(not os.path.isdir("/root/reference"), reason="the reference is only present in the build container")
def test_golden_vectors_are_reproducible(tmp_path, oracle_built):
    """Re-run the reference through the harness and compare with the committed fixtures (build container only)."""
    import sys
    sys.path.insert(0, gc.GOLDEN)
    import ref_harness as rh
    g = gc.load("fit_ref_f32")
    pack = hand_pack.load_pack("allegro")
    B = len(g["pose0"])
    st = rh.make_fit_state(pack, [g["face_verts"]], [g["surface"]], B, g["pose0"], g["idx0"].astype(np.int64))
    np.testing.assert_allclose(st.energy.detach().numpy(), g["energy0"], rtol=1e-6)
    o = rh.fit_iterate(st, torch.tensor(g["t0_u_switch"]), torch.tensor(g["t0_idx_rand"]).long(), torch.tensor(g["t0_u_accept"]))
    np.testing.assert_allclose(o["new_energy"].numpy(), g["t0_new_energy"], rtol=1e-6)
    np.testing.assert_array_equal(o["accept"].numpy(), g["t0_accept"])
    # the reset fixture: regenerate it whole (reference MalaStar.reset_envs, clip_grad, injected NaN) and compare
    import make_golden as mg
    torch.manual_seed(0)
    mg.HERE = str(tmp_path)
    mg.gen_fit_reset()
    mg.gen_metric_alt()  # (f3) DexgraspSpanMetric / TDGSpanMetric
    mg.gen_fit_alt()     # (f3) AnnealingDexGraspNet, the other energies in the loop, the after-reset gradient rule
    mg.gen_models()      # the reference's own HandModel / ObjectModel methods on pack-filled instances
    for name in ("fit_reset_ref_f64", "metric_alt_ref", "fit_alt_ref", "models_ref"):
        new, old = np.load(tmp_path / (name + ".npz")), gc.load(name)
        assert sorted(new.files) == sorted(old)
        for k in new.files:
            if new[k].dtype.kind == "f":
                np.testing.assert_allclose(new[k], old[k], rtol=1e-6 if new[k].dtype == np.float32 else 1e-9, atol=1e-12,
                                           equal_nan=True, err_msg=k)
            else:
                np.testing.assert_array_equal(new[k], old[k], err_msg=k)


@pytest.mark.skipif(not os.path.isdir("/root/reference"), reason="the reference is only present in the build container")
def test_reference_model_classes_reproduce_the_fixtures(oracle_built):
    """The loop fixtures were generated with oracle.hand.HandOracle / oracle.pipeline.ObjectOracle standing in for the
    reference's HandModel / ObjectModel (whose constructors need trimesh / pytorch3d).  Here the SAME runs are repeated with the
    reference's OWN classes — instances made without constructors and filled from the hand pack
    (ref_harness.reference_hand_model / reference_object_model), so that set_parameters (incl. env_mask), fk, cal_distance,
    self_penetration and ObjectModel.cal_distance are the reference's methods — and must reproduce the committed numbers:
    energies, proposals, accept decisions, poses; gradients to summation-order rounding."""
    import sys
    sys.path.insert(0, gc.GOLDEN)
    import ref_harness as rh
    pack = hand_pack.load_pack("allegro")

    def replay(g, prefix, plan, dtype, z_thr=1.0, nan_at=None, **kw):
        B = len(g["pose0"])
        if dtype == torch.float64:
            torch.set_default_dtype(torch.float64)
        try:
            st = rh.make_fit_state(pack, [g["face_verts"]], [g["surface"]], B, g["pose0"], g["idx0"].astype(np.int64),
                                   dtype=dtype, models="reference", **kw)
            tt = lambda a: torch.tensor(a, dtype=dtype)
            rtol = 1e-6 if dtype == torch.float32 else 1e-8  # fp64: pk and the restatement order a few products differently
            for t, kind in enumerate(plan):
                k = f"{prefix}t{t}_"
                if kind == "nan":
                    st.hand.hand_pose.grad[1, 4] = float("nan")
                    st.hand.hand_pose.grad[2, 11] = 1e4
                reset = None
                if str(kind).startswith("reset"):
                    mask = torch.tensor(g[k + "mask_in"]) if k + "mask_in" in g else None
                    reset = (tt(g[k + "pose_init"]), torch.tensor(g[k + "idx_init"]), mask)
                o = rh.fit_iterate(st, tt(g[k + "u_switch"]), torch.tensor(g[k + "idx_rand"]).long(), tt(g[k + "u_accept"]),
                                   reset=reset, z_score_threshold=z_thr)
                np.testing.assert_array_equal(o["idx_prop"].numpy(), g[k + "idx_prop"], err_msg=k)
                np.testing.assert_allclose(o["proposal"].numpy(), g[k + "proposal"], rtol=rtol, atol=1e-7, err_msg=k)
                np.testing.assert_allclose(o["new_energy"].numpy(), g[k + "new_energy"], rtol=10 * rtol, err_msg=k)
                gw = g[k + "new_grad"]
                np.testing.assert_allclose(o["new_grad"].numpy(), gw, rtol=0, atol=1e-5 * np.abs(gw).max(), err_msg=k)
                np.testing.assert_array_equal(o["accept"].numpy(), g[k + "accept"], err_msg=k)
                np.testing.assert_allclose(o["pose"].numpy(), g[k + "pose"], rtol=rtol, atol=1e-7, err_msg=k)
                np.testing.assert_array_equal(o["idx"].numpy(), g[k + "idx"], err_msg=k)
        finally:
            torch.set_default_dtype(torch.float32)

    replay(gc.load("fit_ref_f32"), "", ["plain"] * 4, torch.float32)
    g = gc.load("fit_reset_ref_f64")
    replay(g, "", [str(s) for s in g["plan"]], torch.float64, z_thr=0.3, clip_grad=True)
    g = gc.load("fit_alt_ref")
    replay(g, "dexanneal_", [str(s) for s in g["dexanneal_plan"]], torch.float32, **gc.ALT_RUNS["dexanneal"])
    replay(g, "malareset_", [str(s) for s in g["malareset_plan"]], torch.float64)
