"""Checks against tests/golden/*.npz — outputs of the reference's own code (tests/golden/make_golden.py, ref_harness.py).
Each check takes (lib, be): the emulator library + NumpyBackend on the CPU, libgraspqp_b200.so + TorchBackend on the GPU."""
import ctypes
import os

import numpy as np
import torch

from graspqp_b200 import _lib, engine, hand_pack
import tol

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load(name):
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    return {k: z[k] for k in z.files}


def fc_params(**kw):
    p = _lib.GqpFcParams(friction=0.2, torque_weight=5.0, lower=1.0, upper=21.0, ridge=1e-4, svd_gain=0.1,
                         values_gain=2.0, values_offset=1e-2, eps=5e-2, max_iter=12, w_fc=1.0, w_dis=0.0)
    for k, v in kw.items():
        setattr(p, k, v)
    return p


def check_sdf(lib, be):
    """K1 bit-exact against the reference kernel body on TorchSDF's own test models."""
    g = load("sdf_ref")
    for key in ("cube", "sphere_42"):
        fv, pts = g[key + "_face_verts"], g[key + "_points"]
        P = len(pts)
        dp, df = be.from_host(pts, np.float32), be.from_host(fv.reshape(-1, 9), np.float32)
        d2, sg = be.empty((P,), np.float32), be.empty((P,), np.int32)
        nr, cl = be.empty((P, 3), np.float32), be.empty((P, 3), np.float32)
        _lib.check(lib.gqp_sdf_forward(be.ptr(dp), be.ptr(df), P, len(fv), be.ptr(d2), be.ptr(sg), be.ptr(nr),
                                       be.ptr(cl), be.stream), lib)
        for got, want in ((d2, "_dist2"), (sg, "_sign"), (nr, "_normals"), (cl, "_clst")):
            np.testing.assert_array_equal(be.to_host(got), g[key + want], err_msg=key + want)
        # BVH form: same winner
        m = ctypes.c_void_p()
        fvh = np.ascontiguousarray(fv.reshape(-1, 9), np.float32)
        _lib.check(lib.gqp_mesh_create(fvh.ctypes.data, len(fvh), ctypes.byref(m)), lib)
        _lib.check(lib.gqp_mesh_sdf_forward(m, be.ptr(dp), P, be.ptr(d2), be.ptr(sg), be.ptr(nr), be.ptr(cl), None,
                                            be.stream), lib)
        for got, want in ((d2, "_dist2"), (sg, "_sign"), (nr, "_normals"), (cl, "_clst")):
            np.testing.assert_array_equal(be.to_host(got), g[key + want], err_msg="bvh " + key + want)
        lib.gqp_mesh_destroy(m)


MORE_HANDS = ("robotiq3", "ability_hand")  # fk_more_ref.npz


def check_fk(lib, be, hands=("allegro", "shadow_hand", "panda", "robotiq2")):
    """H2/H3 against pytorch_kinematics' Chain.forward_kinematics and Transform3d.transform_points/normals (for the ability
    hand: fed by the reference's own get_all_joint_angles, hands/ability_hand.py:10-29)."""
    g = load("fk_ref")
    g.update(load("fk_more_ref"))
    for name in hands:
        k = "" if name == "allegro" else name + "_"
        pack = hand_pack.load_pack(name)
        hand = hand_pack.create_hand(lib, pack)
        theta = g[k + "theta"]
        B, L, C = len(theta), len(pack["mesh_frame"]), len(pack["cand"])
        pose = np.zeros((B, 9 + theta.shape[1]), np.float32)
        pose[:, 3], pose[:, 7] = 1.0, 1.0  # identity global rotation (ortho6d), zero translation
        pose[:, 9:] = theta
        dpose, tf = be.from_host(pose, np.float32), be.empty((B, L, 12), np.float32)
        _lib.check(lib.gqp_fk_forward(hand, be.ptr(dpose), B, be.ptr(tf), be.stream), lib)
        np.testing.assert_allclose(be.to_host(tf).reshape(B, L, 3, 4), g[k + "link_matrices"][:, :, :3, :], atol=2e-6,
                                   err_msg=name)
        # candidates in the reference's concatenation order = ascending link, then table order
        order = np.argsort(pack["cand_link"], kind="stable").astype(np.int32)
        idx = np.tile(order, (B, 1))
        cp, cn = be.empty((B, C, 3), np.float32), be.empty((B, C, 3), np.float32)
        d_idx = be.from_host(idx, np.int32)  # kept alive across the call
        _lib.check(lib.gqp_contacts_forward(hand, be.ptr(tf), be.ptr(d_idx), B, C, be.ptr(cp), be.ptr(cn), be.stream), lib)
        np.testing.assert_allclose(be.to_host(cp), g[k + "cand_points"], atol=2e-6, err_msg=name)
        np.testing.assert_allclose(be.to_host(cn), g[k + "cand_normals"], atol=5e-6, err_msg=name)
        lib.gqp_hand_destroy(hand)


def check_metric(lib, be):
    """Q1-Q4 against GraspSpanMetricFactory.create(GRASPQP) (span.py, registry.py, qp_solver.py; qpth = oracle.qp)."""
    g = load("metric_ref")
    for n_c in (12, 16, 4):
        cp, nrm, cog = g[f"nc{n_c}_cpts"], g[f"nc{n_c}_nrm"], g[f"nc{n_c}_cog"]
        B = len(cp)
        P = fc_params()
        d = lambda a: be.from_host(a, np.float32)
        dcp, dn, dcog, zeros = d(cp), d(nrm), d(cog), d(np.zeros((B, n_c)))
        e, e_dis, lam = be.empty((B,), np.float32), be.empty((B,), np.float32), be.empty((B, n_c), np.float32)
        gcp, gcn = be.empty((B, n_c, 3), np.float32), be.empty((B, n_c, 3), np.float32)
        ws = be.empty(((int(lib.gqp_fc_workspace_bytes(B, n_c, P.max_iter)) + 7) // 8,), np.float64)
        p = be.ptr
        # cog per candidate (energy.py:35-42): B "objects" of one candidate
        _lib.check(lib.gqp_fc_solve(ctypes.byref(P), p(dcp), p(dn), p(dcog), B, 1, n_c, p(ws), be.stream), lib)
        _lib.check(lib.gqp_fc_finish(ctypes.byref(P), p(dcp), p(dn), p(zeros), p(dn), None, p(dcog), B, 1, n_c, p(ws), 1,
                                     p(e), p(e_dis), p(lam), p(gcp), p(gcn), None, be.stream), lib)
        want_g = g[f"nc{n_c}_f64_grad_cpts"]
        # tolerances of the north star: energies / gradients 1e-4 relative (fp32 inputs, fp64 reference run)
        np.testing.assert_allclose(be.to_host(e), g[f"nc{n_c}_f64_E_fc"], rtol=1e-4, err_msg=f"E_fc n_c={n_c}")
        np.testing.assert_allclose(be.to_host(lam), g[f"nc{n_c}_f64_x"], rtol=1e-4, atol=1e-4)
        np.testing.assert_allclose(be.to_host(gcp), want_g, rtol=1e-4, atol=1e-4 * np.abs(want_g).max())
        # for the record: the reference's own fp32 run (fp32 LU of the 2n x 2n Schur complement) is only within ~3e-3
        # of its fp64 run on these inputs — the fp64 run is the tighter target, and the one the kernel is held to
        np.testing.assert_allclose(g[f"nc{n_c}_f32_E_fc"], g[f"nc{n_c}_f64_E_fc"], rtol=1e-2)


def check_fit(lib, be, T=4):
    """The whole iteration against scripts/fit.py:381-458 run with the reference MalaStar / calculate_energy / metric
    (fp64 torch on the CPU: the rounding-free run of the reference code) and the same injected noise.  Tolerances (tests/tol.py): energies and gradients 1e-4 relative
    (gradients per pose column, relative to the column's largest entry), accept decisions and contact indices bit-exact."""
    g = load("fit_ref_f64")
    pack = hand_pack.load_pack("allegro")
    B = len(g["pose0"])
    eng = engine.FitEngine(lib, be, pack, [g["face_verts"]], g["surface"][None], B, g["pose0"], g["idx0"])
    H = be.to_host
    np.testing.assert_allclose(H(eng.energy), g["energy0"], rtol=1e-4)
    for k, name in enumerate(engine.TERMS):
        np.testing.assert_allclose(H(eng.losses)[k], g["loss0_" + name], rtol=1e-4, atol=1e-6, err_msg=name)
    tol.assert_grad(eng.first_grad, g["first_grad"], "golden fit: first gradient")
    ex = lambda b: tol.normal_unstable(H(eng.cpts)[b], g["face_verts"])
    n_acc = 0
    for t in range(T):
        eng.step(be.from_host(g[f"t{t}_u_switch"], np.float32), be.from_host(g[f"t{t}_idx_rand"], np.int32),
                 be.from_host(g[f"t{t}_u_accept"], np.float32))
        np.testing.assert_array_equal(H(eng.idx_new), g[f"t{t}_idx_prop"])
        np.testing.assert_allclose(H(eng.pose_new), g[f"t{t}_proposal"], rtol=1e-4, atol=1e-5)
        np.testing.assert_allclose(H(eng.step_size_out), g[f"t{t}_s"], rtol=1e-6)
        np.testing.assert_allclose(H(eng.z), g[f"t{t}_z"], rtol=1e-3, atol=1e-4)
        np.testing.assert_allclose(H(eng.new_energy), g[f"t{t}_new_energy"], rtol=1e-4)
        for k, name in enumerate(engine.TERMS):
            np.testing.assert_allclose(H(eng.losses_new)[k], g[f"t{t}_loss_{name}"], rtol=1e-4, atol=1e-6, err_msg=name)
        tol.assert_grad(H(eng.grad_new), g[f"t{t}_new_grad"], f"golden fit: step {t} gradient", excuse=ex)
        np.testing.assert_allclose(H(eng.temperature), g[f"t{t}_temperature"], rtol=1e-5)
        np.testing.assert_array_equal(H(eng.accept).astype(bool), g[f"t{t}_accept"])
        np.testing.assert_allclose(H(eng.energy), g[f"t{t}_energy"], rtol=1e-4)
        np.testing.assert_allclose(H(eng.pose), g[f"t{t}_pose"], rtol=1e-4, atol=1e-5)
        np.testing.assert_array_equal(H(eng.idx), g[f"t{t}_idx"])
        n_acc += int(g[f"t{t}_accept"].sum())
    eng.close()
    return n_acc


def check_fit_reset(lib, be):
    """M3 (MalaStar.reset_envs, fit.py:408-422), clip_grad with NaN / huge gradient entries (optimizer.py:211-213) against
    the reference's own MalaStar: tests/golden/fit_reset_ref_f64.npz (make_golden.gen_fit_reset).  -> resets seen"""
    g = load("fit_reset_ref_f64")
    pack = hand_pack.load_pack("allegro")
    B = len(g["pose0"])
    eng = engine.FitEngine(lib, be, pack, [g["face_verts"]], g["surface"][None], B, g["pose0"], g["idx0"], clip_grad=True)
    H = be.to_host
    ex = lambda b: tol.normal_unstable(H(eng.cpts)[b], g["face_verts"])
    n_reset = 0
    for t, kind in enumerate(g["plan"]):
        if kind == "nan":  # the same poisoned gradient entries the golden run saw
            gr = H(eng.grad)
            gr[1, 4], gr[2, 11] = np.nan, 1e4
            be.copy_(eng.grad, be.from_host(gr, np.float32))
        kw = {}
        if str(kind).startswith("reset"):
            kw["reset_init"] = (be.from_host(g[f"t{t}_pose_init"], np.float32), be.from_host(g[f"t{t}_idx_init"], np.int32))
            kw["z_score_threshold"] = 0.3
            if f"t{t}_mask_in" in g:
                kw["reset_mask"] = be.from_host(g[f"t{t}_mask_in"].astype(np.uint8), np.uint8)
        eng.step(be.from_host(g[f"t{t}_u_switch"], np.float32), be.from_host(g[f"t{t}_idx_rand"], np.int32),
                 be.from_host(g[f"t{t}_u_accept"], np.float32), **kw)
        if kw:
            np.testing.assert_array_equal(H(eng.reset_mask).astype(bool), g[f"t{t}_reset_mask"], err_msg=f"step {t}")
            assert int(H(eng.reset_count)[0]) == int(g[f"t{t}_reset_mask"].sum())
            n_reset += int(g[f"t{t}_reset_mask"].sum())
        np.testing.assert_array_equal(H(eng.idx_new), g[f"t{t}_idx_prop"], err_msg=f"step {t}")
        np.testing.assert_allclose(H(eng.pose_new), g[f"t{t}_proposal"], rtol=1e-4, atol=1e-5, err_msg=f"step {t}")
        np.testing.assert_allclose(H(eng.z), g[f"t{t}_z"], rtol=1e-3, atol=1e-4)
        tol.assert_energy(H(eng.new_energy), g[f"t{t}_new_energy"], f"golden reset run: step {t} ({kind}) new energy", excuse=ex)
        for k, name in enumerate(engine.TERMS):
            np.testing.assert_allclose(H(eng.losses_new)[k], g[f"t{t}_loss_{name}"], rtol=1e-4, atol=1e-6, err_msg=name)
        tol.assert_grad(H(eng.grad_new), g[f"t{t}_new_grad"], f"golden reset run: step {t} ({kind}) gradient", excuse=ex)
        np.testing.assert_allclose(H(eng.temperature), g[f"t{t}_temperature"], rtol=1e-4)  # through z (rtol 1e-3)
        np.testing.assert_array_equal(H(eng.accept).astype(bool), g[f"t{t}_accept"], err_msg=f"step {t}")
        np.testing.assert_allclose(H(eng.energy), g[f"t{t}_energy"], rtol=1e-4)
        np.testing.assert_allclose(H(eng.pose), g[f"t{t}_pose"], rtol=1e-4, atol=1e-5)
        np.testing.assert_array_equal(H(eng.idx), g[f"t{t}_idx"])
        np.testing.assert_array_equal(H(eng.step_count), g[f"t{t}_step"], err_msg=f"step counters after iteration {t}")
        np.testing.assert_allclose(H(eng.ema), g[f"t{t}_ema"], rtol=1e-4, atol=1e-9, err_msg=f"ema after iteration {t}")
    eng.close()
    return n_reset


def check_metric_alt(lib, be):
    """(f3) gqp_span_energy against the reference's DexgraspSpanMetric (torque weight 0 — its path's value — and 5) and
    TDGSpanMetric outputs and autograd gradients: tests/golden/metric_alt_ref.npz (make_golden.gen_metric_alt, fp32 runs of
    the reference, which holds fp32 constants).  E_dis and its two gradients against the closed form of energy.py:25-28."""
    import ctypes
    from graspqp_b200 import _lib
    g = load("metric_alt_ref")
    p = be.ptr
    for n_c in (12, 4):
        k = f"nc{n_c}_"
        cp, nrm, cog = g[k + "cpts"], g[k + "nrm"], g[k + "cog"]
        B = len(cp)
        rng = np.random.default_rng(1)
        nh = rng.normal(size=(B, n_c, 3)).astype(np.float32)
        dist = (rng.normal(size=(B, n_c)) * 0.01).astype(np.float32)
        dg = rng.normal(size=(B, n_c, 3)).astype(np.float32)
        d_cp, d_nrm, d_cog, d_nh, d_dist, d_dg = [be.from_host(a, np.float32) for a in (cp, nrm, cog, nh, dist, dg)]
        d_dirs = be.from_host(g[k + "tdg_dirs"], np.float32)
        e, ed = be.empty((B,), np.float32), be.empty((B,), np.float32)
        gc_, gn = be.empty((B, n_c, 3), np.float32), be.empty((B, n_c, 3), np.float32)
        for kind, tw, key in ((_lib.SPAN_DEXGRASP, 0.0, "dex_tw0"), (_lib.SPAN_DEXGRASP, 5.0, "dex_tw5"),
                              (_lib.SPAN_TDG, 0.0, "tdg")):
            P = _lib.GqpSpanParams(kind=kind, torque_weight=tw, friction=0.2, obb_length=0.2, scale=100.0,
                                   enable_density=1, w_fc=1.0, w_dis=0.0)
            _lib.check(lib.gqp_span_energy(ctypes.byref(P), p(d_cp), p(d_nh), p(d_dist), p(d_nrm), p(d_dg), p(d_cog),
                                           p(d_dirs), len(g[k + "tdg_dirs"]), B, 1, n_c, p(e), p(ed), p(gc_), p(gn),
                                           be.stream), lib)
            tol.assert_energy(be.to_host(e), g[k + key + "_E"], f"golden {key} n_c={n_c} energy", max_excused=0)
            want = g[k + key + "_grad"]
            err = np.abs(be.to_host(gc_) - want).max() / max(np.abs(want).max(), 1e-30)
            print(f"[grad] golden {key} n_c={n_c}: d/d contact points, error / largest entry {err:.2e}; tolerance 1e-04")
            assert err < 1e-4
        P = _lib.GqpSpanParams(kind=_lib.SPAN_DEXGRASP, torque_weight=0.0, friction=0.2, obb_length=0.2, scale=100.0,
                               enable_density=1, w_fc=0.0, w_dis=100.0)
        _lib.check(lib.gqp_span_energy(ctypes.byref(P), p(d_cp), p(d_nh), p(d_dist), p(d_nrm), p(d_dg), p(d_cog), None, 0,
                                       B, 1, n_c, p(e), p(ed), p(gc_), p(gn), be.stream), lib)
        ea = np.exp(1 + (nrm.astype(np.float64) * nh).sum(-1))
        np.testing.assert_allclose(be.to_host(ed), (ea * np.abs(dist)).sum(-1), rtol=1e-5)
        np.testing.assert_allclose(be.to_host(gn), 100 * (ea * np.abs(dist))[..., None] * nrm, rtol=1e-4, atol=1e-5)
        np.testing.assert_allclose(be.to_host(gc_), 100 * (ea * np.sign(dist))[..., None] * dg, rtol=1e-4, atol=1e-3)
    # argument checks of the entry point
    P = _lib.GqpSpanParams(kind=7)
    assert lib.gqp_span_energy(ctypes.byref(P), p(d_cp), p(d_nh), p(d_dist), p(d_nrm), None, p(d_cog), None, 0, B, 1, n_c,
                               p(e), p(ed), None, None, be.stream) != 0
    P = _lib.GqpSpanParams(kind=_lib.SPAN_TDG, friction=0.2, obb_length=0.2, scale=100.0)
    assert lib.gqp_span_energy(ctypes.byref(P), p(d_cp), p(d_nh), p(d_dist), p(d_nrm), None, p(d_cog), None, 0, B, 1, n_c,
                               p(e), p(ed), None, None, be.stream) != 0  # tdg without its direction table
    assert lib.gqp_span_energy(ctypes.byref(P), p(d_cp), p(d_nh), p(d_dist), p(d_nrm), None, p(d_cog), p(d_dirs), 1000, B, 1,
                               33, p(e), p(ed), None, None, be.stream) != 0  # n_c beyond 32


ALT_RUNS = {
    "dexanneal": dict(energy_type="dexgrasp", optimizer="dexgraspnet", switch_possibility=0.5, stepsize_period=2,
                      annealing_period=3),
    "tdgmala": dict(energy_type="tdg"),
    "malareset": dict(),
    "priorwall": dict(weights=dict(E_prior=5.0, E_wall=50.0)),
}


def check_fit_alt(lib, be, runs=("dexanneal", "tdgmala", "malareset", "priorwall")):
    """(f3) the reference's OTHER optimiser / energies and the gradient a rejected row keeps after a reset iteration, against
    the reference's own AnnealingDexGraspNet, DexgraspSpanMetric, TDGSpanMetric and MalaStar run through scripts/fit.py's loop
    body: tests/golden/fit_alt_ref.npz (make_golden.gen_fit_alt; "dexanneal" is an fp32 run of the reference, the other two
    fp64).  -> number of rejected rows seen in iterations that follow a reset iteration."""
    g = load("fit_alt_ref")
    pack = hand_pack.load_pack("allegro")
    B = len(g["pose0"])
    H = be.to_host
    n_after = 0
    for run in runs:
        kw = dict(ALT_RUNS[run])
        if run == "tdgmala":
            kw["tdg_directions"] = g["tdg_dirs"]
        if run == "priorwall":
            kw["hand_surface"] = (g["hand_surf"], g["hand_surf_link"])
        eng = engine.FitEngine(lib, be, pack, [g["face_verts"]], g["surface"][None], B, g["pose0"], g["idx0"], **kw)
        tol.assert_energy(H(eng.energy), g[run + "_energy0"], f"golden {run}: first energy")
        tol.assert_grad(eng.first_grad, g[run + "_first_grad"], f"golden {run}: first gradient")
        ex = lambda b: tol.normal_unstable(H(eng.cpts)[b], g["face_verts"])
        prev_reset = False
        for t, kind in enumerate(g[run + "_plan"]):
            k0 = f"{run}_t{t}_"
            kw = {}
            if kind == "reset_mask":
                kw["reset_init"] = (be.from_host(g[k0 + "pose_init"], np.float32), be.from_host(g[k0 + "idx_init"], np.int32))
                kw["reset_mask"] = be.from_host(g[k0 + "mask_in"].astype(np.uint8), np.uint8)
            eng.step(be.from_host(g[k0 + "u_switch"], np.float32), be.from_host(g[k0 + "idx_rand"], np.int32),
                     be.from_host(g[k0 + "u_accept"], np.float32), **kw)
            lbl = f"golden {run}: step {t} ({kind})"
            np.testing.assert_array_equal(H(eng.idx_new), g[k0 + "idx_prop"], err_msg=lbl)
            np.testing.assert_allclose(H(eng.pose_new), g[k0 + "proposal"], rtol=1e-4, atol=1e-5, err_msg=lbl)
            np.testing.assert_allclose(np.broadcast_to(g[k0 + "s"], (B,)) if run == "dexanneal" else g[k0 + "s"],
                                       H(eng.step_size_out)[:1].repeat(B) if run == "dexanneal" else H(eng.step_size_out),
                                       rtol=1e-6, err_msg=lbl)
            tol.assert_energy(H(eng.new_energy), g[k0 + "new_energy"], lbl + " new energy", excuse=ex)
            for k, name in enumerate(eng.terms):
                np.testing.assert_allclose(H(eng.losses_new)[k], g[k0 + "loss_" + name], rtol=1e-4, atol=1e-6, err_msg=name)
            tol.assert_grad(H(eng.grad_new), g[k0 + "new_grad"], lbl + " gradient", excuse=ex)
            np.testing.assert_allclose(H(eng.temperature), np.broadcast_to(g[k0 + "temperature"], (B,)), rtol=1e-4,
                                       err_msg=lbl)
            np.testing.assert_array_equal(H(eng.accept).astype(bool), g[k0 + "accept"], err_msg=lbl)
            np.testing.assert_allclose(H(eng.energy), g[k0 + "energy"], rtol=1e-4, err_msg=lbl)
            np.testing.assert_allclose(H(eng.pose), g[k0 + "pose"], rtol=1e-4, atol=1e-5, err_msg=lbl)
            np.testing.assert_array_equal(H(eng.idx), g[k0 + "idx"], err_msg=lbl)
            if run == "dexanneal":
                assert eng.anneal_step == int(g[k0 + "step"])
                np.testing.assert_allclose(H(eng.ema)[eng.ema_cur], g[k0 + "ema"], rtol=1e-4, atol=1e-9, err_msg=lbl)
            else:
                np.testing.assert_array_equal(H(eng.step_count), g[k0 + "step"], err_msg=lbl)
                # the accumulator is a mean of SQUARED gradients: twice their relative error (tolerance 1e-4)
                np.testing.assert_allclose(H(eng.ema), g[k0 + "ema"], rtol=3e-4, atol=1e-9, err_msg=lbl)
            if prev_reset:
                n_after += int((~g[k0 + "accept"]).sum())
            prev_reset = kind == "reset_mask"
        eng.close()
    return n_after


MODEL_HANDS = ("allegro", "shadow_hand", "panda", "ability_hand", "robotiq3")


def models_case(hand, seed=3):
    """the inputs make_golden.gen_models used (same construction; the fixture stores only the reference's outputs)"""
    from graspqp_b200 import synth
    pack = hand_pack.load_pack(hand)
    v, f = synth.make_mesh("sphere", 400, seed)
    B = 4
    pose, idx = synth.random_grasp_init(pack, [v], B, 4, seed)
    pose[:, :3] *= 0.6
    lo, hi = pack["joints_lower"], pack["joints_upper"]
    pose[B // 2:, 9:] = (lo + 0.92 * (hi - lo))[None]
    sp = synth.sample_surface(v, f, 96, seed=seed)
    return pack, synth.face_vertices(v, f), pose.astype(np.float32), idx, sp.astype(np.float32), hand_pack.surface_points(pack, 64)


def check_models(lib, be, hands=MODEL_HANDS):
    """H4 / H5 / O1 kernels against the reference's OWN HandModel.cal_distance, self_penetration and ObjectModel.cal_distance
    (tests/golden/models_ref.npz, make_golden.gen_models: the reference methods run on instances filled from the hand packs).
    Distances: the kernels answer max over links of the same per-link TorchSDF arithmetic -> 1e-6 absolute (the link transforms
    come from the fp32 FK kernel, a few ulps from pytorch_kinematics'); E_spen 1e-5; object distance / normals of the given
    points to 1 ulp (same points in, K1 arithmetic bit-exact, one square root on top)."""
    g = load("models_ref")
    p = be.ptr
    for hand in hands:
        pack, fv, pose, idx, sp, surf = models_case(hand)
        k = hand + "_"
        B, n_pts = len(pose), len(sp)
        h = hand_pack.create_hand(lib, pack)
        L = len(pack["mesh_frame"])
        d_pose, tf = be.from_host(pose, np.float32), be.empty((B, L, 12), np.float32)
        _lib.check(lib.gqp_fk_forward(h, p(d_pose), B, p(tf), be.stream), lib)
        # H4, dist-returning path: per-sample signed depth (the surface samples are per OBJECT: n_obj = 1, batch_each = B)
        x = be.from_host(sp[None], np.float32)
        dist, arg = be.empty((B, n_pts), np.float32), be.empty((B, n_pts), np.int32)
        dg = be.empty((B, n_pts, 3), np.float32)
        sc = be.empty((int(lib.gqp_hand_distance_scratch_floats(1, B, n_pts, 0)),), np.float32)
        _lib.check(lib.gqp_hand_distance_full(h, p(tf), p(x), 1, B, n_pts, 0, p(dist), p(arg), p(dg), p(sc), be.stream), lib)
        want = g[k + "hand_distance"]
        err = np.abs(be.to_host(dist) - want).max()
        print(f"[models] {hand}: HandModel.cal_distance max |error| {err:.2e} m (max depth {want.max():.4f})")
        assert err < 1e-6
        # H5
        e_spen, e_j = be.empty((B,), np.float32), be.empty((B,), np.float32)
        _lib.check(lib.gqp_self_pen_joints(h, p(tf), p(d_pose), B, p(e_spen), p(e_j), 1.0, 1.0, None, None, be.stream), lib)
        np.testing.assert_allclose(be.to_host(e_spen), g[k + "e_spen"], rtol=1e-4, atol=1e-6, err_msg=hand)
        # O1 at the stored points
        pts = g[k + "obj_points"]
        m = ctypes.c_void_p()
        fv9 = np.ascontiguousarray(fv, np.float32).reshape(-1, 9)
        _lib.check(lib.gqp_mesh_create(fv9.ctypes.data, len(fv9), ctypes.byref(m)), lib)
        arr = (ctypes.c_void_p * 1)(m.value)
        objset = ctypes.c_void_p()
        _lib.check(lib.gqp_objset_create(ctypes.cast(arr, ctypes.c_void_p), 1, ctypes.byref(objset)), lib)
        n = pts.shape[1]
        d_pts = be.from_host(pts, np.float32)
        od, on = be.empty((B, n), np.float32), be.empty((B, n, 3), np.float32)
        oc, osg = be.empty((B, n, 3), np.float32), be.empty((B, n), np.int32)
        _lib.check(lib.gqp_object_distance_forward(objset, B, n, p(d_pts), 1.0, p(od), p(on), p(oc), p(osg), be.stream), lib)
        # d2 is bit-exact (K1); torch's CPU sqrt, which made the fixture, is not always the correctly rounded one: 1 ulp
        np.testing.assert_allclose(be.to_host(od), g[k + "obj_distance"], rtol=2e-7, atol=0, err_msg=hand)
        np.testing.assert_allclose(be.to_host(on), g[k + "obj_normals"], rtol=0, atol=2e-7, err_msg=hand)
        ones, ograd = be.from_host(np.sign(g[k + "obj_distance"]).astype(np.float32), np.float32), be.empty((B, n, 3), np.float32)
        _lib.check(lib.gqp_object_distance_backward(p(ones), p(d_pts), p(oc), p(osg), B * n, 1.0, p(ograd), be.stream), lib)
        np.testing.assert_allclose(be.to_host(ograd), g[k + "obj_grad"], rtol=1e-5, atol=1e-6, err_msg=hand)
        lib.gqp_objset_destroy(objset)
        lib.gqp_mesh_destroy(m)
        lib.gqp_hand_destroy(h)
