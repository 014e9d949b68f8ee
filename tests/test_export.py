"""(f2) export path (graspqp_b200/export.py) vs restatements of scripts/fit.py:224-300 and
HandModel.get_req_joint_velocities (hand_model.py:1155-1218), and the contact Jacobian against pytorch_kinematics' own
vectorised Jacobian (pk/chain.py:197-252) run in this container."""
import os
import sys

import numpy as np
import pytest
import torch

from graspqp_b200 import engine, export, hand_pack, synth
from oracle.hand import HandOracle


def _engine(emu, hand="allegro", B=3, n_c=4, seed=2, be=None):
    pack = hand_pack.load_pack(hand)
    v, f = synth.make_mesh("sphere", 200, seed)
    fv, sp = synth.face_vertices(v, f), synth.sample_surface(v, f, 32, seed=seed)
    pose, idx = synth.random_grasp_init(pack, [v], B, n_c, seed)
    pose[:, :3] *= 0.8
    return pack, engine.FitEngine(emu, be or engine.NumpyBackend(), pack, [fv], sp[None], B, pose, idx.astype(np.int32)), fv


def test_quaternion_matches_scipy():
    from scipy.spatial.transform import Rotation
    R = Rotation.random(200, random_state=0)
    q = export.rotmat_to_quat_wxyz(torch.tensor(R.as_matrix(), dtype=torch.float64)).numpy()
    want = R.as_quat()[:, [3, 0, 1, 2]]
    want = want * np.sign(want[:, :1])
    np.testing.assert_allclose(q, want, atol=1e-12)


@pytest.mark.parametrize("hand", ["allegro", "panda", "ability_hand", "robotiq2"])
def test_contact_jacobian_and_joint_velocities(emu, hand):
    check_jacobian_and_velocities(emu, None, hand)


def check_jacobian_and_velocities(lib, be, hand, B=3, n_c=4):
    pack, eng, fv = _engine(lib, hand, B=B, n_c=n_c, be=be)
    H = eng.be.to_host
    ex = export.Exporter(eng)
    J = H(ex.contact_jacobian(eng.idx))  # (Bt, n_c, 3, dof), world frame
    # oracle: autograd through the restated HandModel (pk FK incl. the +-0.01 rad perturbation)
    hp = torch.tensor(H(eng.pose), dtype=torch.float64)
    idx = torch.tensor(H(eng.idx)).long()
    if "mlp_w1" in pack:
        # Robotiq 2F-140: the reference's jacobian_fnc (hands/robotiq2.py:17-20) = column(finger_joint) - column(right outer
        # knuckle) of the Jacobian w.r.t. the EXPANDED joints, not the derivative through the coupling network
        o = HandOracle(hand_pack.expanded_pack(pack), torch.float64)
        th0 = torch.tensor(hand_pack.joint_vector(pack, H(eng.pose)[:, 9:]), dtype=torch.float64)
    else:
        o = HandOracle(pack, torch.float64)
        th0 = hp[:, 9:].clone()

    def cpts(theta):
        o.set_parameters(torch.cat([hp[:, :9], theta], 1), idx)
        return o.contact_points
    Jw = torch.autograd.functional.jacobian(lambda th: cpts(th).sum(0), th0)  # (n_c,3,Bt,dof)
    Jw = Jw.permute(2, 0, 1, 3)
    if "mlp_w1" in pack:
        Jw = sum(wgt * Jw[..., col:col + 1] for col, wgt in hand_pack.reference_jacobian_columns(pack))
        o = HandOracle(pack, torch.float64)
        o.set_parameters(hp, idx)  # the global rotation below
    Jw = Jw.numpy()
    np.testing.assert_allclose(J, Jw, atol=2e-6)
    # get_req_joint_velocities restated in the reference's own frame conventions (hand frame + rotated directions)
    Rg = o.global_rotation.detach()
    v = torch.tensor(np.random.default_rng(0).normal(size=(eng.Bt, eng.n_c, 3)))
    Jh = (Rg.mT.unsqueeze(1) @ torch.tensor(Jw))                       # hand-frame Jacobian, as hand_model.py:1190-1196
    vh = (Rg.mT.unsqueeze(1) @ v.unsqueeze(-1)).squeeze(-1)            # :1164
    want = (export.pinv(Jh.flatten(1, 2)) @ vh.flatten(1, 2).unsqueeze(-1)).squeeze(-1)
    theta, res, ee = ex.get_req_joint_velocities(eng.be.from_host(v.float().numpy(), np.float32), eng.idx, return_ee_vel=True)
    np.testing.assert_allclose(theta.cpu().numpy(), want.numpy(), rtol=2e-3, atol=2e-4 * np.abs(want.numpy()).max())
    ex.close()
    eng.close()


def test_jacobian_linear_part_is_pk_jacobian():
    """pk's Chain.jacobian (vectorised, what HandModel.jacobian calls) == d(link position)/d(theta): the premise of
    building the contact Jacobian from the FK backward kernels"""
    if not os.path.isdir("/root/reference"):
        pytest.skip("reference only in the build container")
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
    import ref_harness as rh
    pack = hand_pack.load_pack("allegro")
    chain, frames = rh.pk_chain_from_pack(pack, torch.float32)
    th = torch.tensor(np.random.default_rng(0).uniform(0.1, 0.6, (2, 16)), dtype=torch.float32)
    J = chain.jacobian(th)

    def pos(t):
        res = chain.forward_kinematics(t)
        return torch.stack([res[f"l{i}"].get_matrix()[:, :3, 3].expand(2, 3) for i in range(len(frames))], 1)
    Jt = torch.autograd.functional.jacobian(lambda t: pos(t).sum(0), th).permute(2, 0, 1, 3)
    np.testing.assert_allclose(J[:, :, :3, :].numpy(), Jt.numpy(), atol=1e-6)


def test_ability_hand_coupled_jacobian_is_the_references():
    """hands/ability_hand.py:36-42 (calculate_jacobian: active columns + 1.05851325 x the mimic joints' columns of pk's
    Jacobian) == d(link position)/d(the 6 parameters) through the pack's frame_gain — what the FK backward kernels compute."""
    if not os.path.isdir("/root/reference"):
        pytest.skip("reference only in the build container")
    import importlib.util
    import types
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
    import ref_harness as rh
    rh.load()
    core = sys.modules["graspqp.core"]
    had = hasattr(core, "HandModel")
    if not had:
        core.HandModel = type("HandModel", (), {})  # the module only names it in an annotation; the class needs trimesh
    spec = importlib.util.spec_from_file_location(
        "ref_ability_hand_j", os.path.join(rh.REF, "graspqp", "src", "graspqp", "hands", "ability_hand.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    if not had:
        del core.HandModel
    pack = hand_pack.load_pack("ability_hand")
    chain, frames = rh.pk_chain_from_pack(pack, torch.float32)  # pk's vectorised Jacobian mixes in float32 constants
    th = torch.tensor(np.random.default_rng(1).uniform(0.2, 0.9, (2, 6)), dtype=torch.float32)
    Jref = mod.calculate_jacobian(th, types.SimpleNamespace(chain=chain))  # (2, n_frames, 6, 6 parameters)
    o = HandOracle(pack, torch.float32)

    def pos(t):
        return torch.stack([m[:, :3, 3] for m in o.fk(t)], 1)  # (2, n_frames, 3): frame_gain applied inside
    Jt = torch.autograd.functional.jacobian(lambda t: pos(t).sum(0), th).permute(2, 0, 1, 3)
    np.testing.assert_allclose(Jref[:, :, :3, :].numpy(), Jt.numpy(), atol=2e-6)


def test_robotiq2_jacobian_is_the_references():
    """hands/robotiq2.py:17-20 (calculate_jacobian over pk's Jacobian at the joints its own robotiq2f140_fk expands) ==
    the column combination export.contact_jacobian applies (hand_pack.reference_jacobian_columns) to the Jacobian w.r.t. the
    expanded joints"""
    if not os.path.isdir("/root/reference"):
        pytest.skip("reference only in the build container")
    import importlib.util
    import types
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
    import ref_harness as rh
    rh.load()
    core = sys.modules["graspqp.core"]
    had = hasattr(core, "HandModel")
    if not had:
        core.HandModel = type("HandModel", (), {})
    try:
        spec = importlib.util.spec_from_file_location(
            "ref_robotiq2_j", os.path.join(rh.REF, "graspqp", "src", "graspqp", "hands", "robotiq2.py"))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
    finally:
        if not had:
            del core.HandModel
    pack = hand_pack.load_pack("robotiq2")
    chain, frames = rh.pk_chain_from_pack(pack, torch.float32)
    movable = [f for f in range(len(pack["frame_jtype"])) if int(pack["frame_jtype"][f]) != 0]
    names = [str(pack["frame_joint"][f]) for f in movable]  # HandModel.joints_names: DFS over non-fixed joints
    th = torch.tensor([[0.1], [0.45], [0.9]], dtype=torch.float32)  # the last one beyond the clamp
    with torch.no_grad():
        Jref = mod.calculate_jacobian(th, types.SimpleNamespace(chain=chain, joints_names=names))  # (3, n_frames, 6, 1)
    qpack = hand_pack.expanded_pack(pack)
    o = HandOracle(qpack, torch.float32)
    q = torch.tensor(hand_pack.joint_vector(pack, th.numpy()))

    def pos(t):
        return torch.stack([m[:, :3, 3] for m in o.fk(t)], 1)
    Jq = torch.autograd.functional.jacobian(lambda t: pos(t).sum(0), q).permute(2, 0, 1, 3)  # (3, n_frames, 3, 6)
    J = sum(wgt * Jq[..., col:col + 1] for col, wgt in hand_pack.reference_jacobian_columns(pack))
    np.testing.assert_allclose(Jref[:, :, :3, :].numpy(), J.numpy(), atol=2e-6)


def test_export_writes_the_reference_dict(emu, tmp_path):
    pack, eng, fv = _engine(emu, "allegro", B=2, n_c=4)
    ex = export.Exporter(eng)
    paths = ex.export_poses(str(tmp_path), ["obj_a"], "allegro", suffix="_0")
    assert paths[0].endswith(os.path.join("obj_a", "grasp_predictions", "allegro", "4_contacts", "graspqp", "default",
                                          "obj_a_0.dexgrasp.pt"))
    d = torch.load(paths[0])
    assert set(d) == {"values", "parameters", "grasp_velocities", "full_grasp_velocities", "grasp_velocities_off",
                      "contact_idx", "grasp_type", "contact_links"}  # fit.py:253-297
    names = [str(s) for s in pack["joint_names"]]
    assert set(d["parameters"]) == set(names) | {"root_pose"} and d["parameters"]["root_pose"].shape == (2, 7)
    assert set(d["grasp_velocities"]) == set(names) and d["contact_idx"].shape == (2, 4)
    np.testing.assert_allclose(d["parameters"]["root_pose"][:, :3].numpy(), eng.pose[:, :3])
    np.testing.assert_allclose(d["parameters"]["root_pose"][:, 3:].norm(dim=1).numpy(), 1.0, atol=1e-6)
    assert all(torch.isfinite(v).all() for v in d["full_grasp_velocities"].values())
    eng.close()


def test_exported_file_is_read_by_the_references_loader(emu, tmp_path):
    """(f2) round trip: the .dexgrasp.pt written here goes through the reference's own consumer,
    graspqp_isaaclab/utils/data.py:105-165 (get_saved_poses), with the joint order of its Isaac Lab asset configuration
    (graspqp_isaaclab/assets/allegro.py:9-26).  The module's Isaac Lab / matplotlib / transforms3d imports are stubbed (none of
    them is used by that function); the simulator itself is unavailable on this box."""
    ref = "/root/reference/graspqp_isaaclab/src/graspqp_isaaclab"
    if not os.path.isdir(ref):
        pytest.skip("reference only in the build container")
    import ast
    import importlib.util
    import types
    stubs = {}
    for name in ("isaaclab", "isaaclab.sim", "isaaclab.assets", "isaaclab.assets.articulation", "isaaclab.utils",
                 "isaaclab.utils.math", "transforms3d", "transforms3d.euler", "matplotlib"):
        stubs[name] = sys.modules.get(name)
        sys.modules[name] = types.ModuleType(name)
    sys.modules["isaaclab.assets.articulation"].Articulation = object
    sys.modules["isaaclab.utils.math"].quat_from_matrix = None
    sys.modules["transforms3d.euler"].euler2mat = None
    sys.modules["matplotlib"].cm = None
    try:
        spec = importlib.util.spec_from_file_location("ref_isaaclab_data", os.path.join(ref, "utils", "data.py"))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
    finally:
        for name, old in stubs.items():
            if old is None:
                del sys.modules[name]
            else:
                sys.modules[name] = old
    tree = ast.parse(open(os.path.join(ref, "assets", "allegro.py")).read())
    names = next(ast.literal_eval(n.value) for n in tree.body if isinstance(n, ast.Assign)
                 and getattr(n.targets[0], "id", "") == "ALLEGRO_HAND_ACTUATED_JOINT_NAMES")
    pack, eng, fv = _engine(emu, "allegro", B=4, n_c=4)
    assert sorted(names) == sorted(str(s) for s in pack["joint_names"])
    path = export.Exporter(eng).export_poses(str(tmp_path), ["obj_a"], "allegro", suffix="_step_7000")[0]
    art = types.SimpleNamespace(cfg=types.SimpleNamespace(actuated_joints_expr=names))
    hand_poses, joint_positions, velocities, energies = mod.get_saved_poses(path, art, num_grasps=3, energy_th=-1e3)
    order = np.argsort(np.asarray(eng.energy), kind="stable")[:3]  # sorted by energy, best first
    np.testing.assert_allclose(energies.numpy(), np.asarray(eng.energy)[order], rtol=1e-6)
    np.testing.assert_allclose(hand_poses[:, :3].numpy(), np.asarray(eng.pose)[order, :3], rtol=1e-6)
    col = [list(str(s) for s in pack["joint_names"]).index(n) for n in names]  # the asset's joint order
    np.testing.assert_allclose(joint_positions.numpy(), np.asarray(eng.pose)[order][:, 9:][:, col], rtol=1e-6)
    assert velocities.shape == (3, 16) and torch.isfinite(velocities).all()
    eng.close()


@pytest.mark.parametrize("hand", ["allegro", "panda", "robotiq2", "ability_hand"])
def test_joint_velocities_match_the_references_own_method(emu, hand):
    """(f2) Exporter.get_req_joint_velocities / contact_jacobian (FK backward kernels, world frame) vs the reference's OWN
    HandModel.get_req_joint_velocities and get_contact_jacobian (hand_model.py:1079-1218: pk's vectorised link Jacobian, the
    per-hand jacobian_fnc, omega x r per contact, hand frame, damped pinv) run on a constructor-less instance filled from the
    pack (ref_harness.reference_hand_model), which first runs the reference's set_parameters on the engine's state."""
    if not os.path.isdir("/root/reference"):
        pytest.skip("reference only in the build container")
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
    import ref_harness as rh
    HandModel, _ = rh.load_models()
    pack, eng, fv = _engine(emu, hand, B=3, n_c=4)
    H = eng.be.to_host
    h = rh.reference_hand_model(pack)
    idx = torch.tensor(H(eng.idx)).long()
    HandModel.set_parameters(h, torch.tensor(H(eng.pose)), idx)
    v = torch.tensor(np.random.default_rng(0).normal(size=(eng.Bt, eng.n_c, 3)).astype(np.float32))
    with torch.no_grad():
        theta_r, res_r, ee_r = HandModel.get_req_joint_velocities(h, v, idx, return_ee_vel=True)
        J_r = HandModel.get_contact_jacobian(h, idx)  # (Bt, n_c, 3, dof), world frame
    ex = export.Exporter(eng)
    J = H(ex.contact_jacobian(eng.idx))
    np.testing.assert_allclose(J, J_r.numpy(), atol=3e-6)
    theta, res, ee = ex.get_req_joint_velocities(eng.be.from_host(v.numpy(), np.float32), eng.idx, return_ee_vel=True)
    scale = np.abs(theta_r.numpy()).max()
    np.testing.assert_allclose(theta.cpu().numpy(), theta_r.numpy(), rtol=2e-3, atol=2e-4 * scale)
    np.testing.assert_allclose(ee.cpu().numpy(), ee_r.numpy(), rtol=2e-3, atol=2e-4 * np.abs(ee_r.numpy()).max())
    ex.close()
    eng.close()
