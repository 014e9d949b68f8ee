"""The reference-shaped Python API (HandModel / ObjectModel / energy callable / MalaStar, autograd-driven as in
scripts/fit.py:381-458) vs the fused FitEngine and vs the oracle, same injected noise.  -m gpu."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

WEIGHTS = dict(E_dis=100.0, E_fc=1.0, E_pen=100.0, E_spen=10.0, E_joints=1.0)


def _case(B=8, n_pts=256, n_c=12, seed=3):
    from graspqp_b200 import hand_pack, synth
    pack = hand_pack.load_pack("allegro")
    v, f = synth.make_mesh("sphere", 400, seed)
    fv = synth.face_vertices(v, f)
    sp = synth.sample_surface(v, f, n_pts, seed=seed)
    pose, idx = synth.random_grasp_init(pack, [v], B, n_c, seed)
    pose[:, :3] *= 0.7
    return pack, fv, sp, pose, idx


def _total(losses):
    e = 0
    for k, v in losses.items():
        e = e + WEIGHTS[k] * v
    return e


def test_reference_shaped_loop_matches_engine_and_oracle(oracle_built):
    import graspqp_b200 as g
    from graspqp_b200 import _lib, engine
    from oracle.pipeline import FitLoopOracle
    B, n_c, T = 8, 12, 3
    pack, fv, sp, pose, idx = _case(B=B, n_c=n_c)
    dev = "cuda:0"
    hand = g.HandModel("allegro", dev)
    obj = g.ObjectModel(batch_size_each=B, num_samples=len(sp), device=dev)
    obj.initialize(["sphere"], [fv], sp[None])
    energy_fnc = g.GraspSpanMetricFactory.create(g.GraspSpanMetricFactory.MetricType.GRASPQP,
                                                 solver_kwargs={"friction": 0.2, "max_limit": 20.0, "n_cone_vecs": 4})
    hand.set_parameters(torch.tensor(pose, device=dev).requires_grad_(True), torch.tensor(idx, device=dev))
    opt = g.MalaStar(hand, switch_possibility=0.4, device=dev, batch_size=B)
    # scripts/fit.py:381-396
    losses = g.calculate_energy(hand, obj, energy_fnc=energy_fnc, method="gendexgrasp", svd_gain=0.1)
    energy = _total(losses)
    energy.sum().backward()
    first_grad = hand.hand_pose.grad.clone()
    opt.zero_grad()
    energy = energy.detach()

    eng = engine.FitEngine(_lib.lib(), engine.TorchBackend(dev), pack, [fv], sp[None], B, pose, idx.astype(np.int32))
    orc = FitLoopOracle(pack, [fv], [sp], B, pose, idx, dtype=torch.float64)
    np.testing.assert_allclose(energy.cpu().numpy(), orc.energy.numpy(), rtol=1e-4)
    np.testing.assert_allclose(energy.cpu().numpy(), eng.energy.cpu().numpy(), rtol=1e-5)
    g0 = orc.first_grad.numpy()
    np.testing.assert_allclose(first_grad.cpu().numpy(), g0, rtol=1e-4, atol=1e-4 * np.abs(g0).max())

    rng = np.random.default_rng(0)
    for t in range(T):
        us, ir, ua = rng.random((B, n_c), dtype=np.float32), rng.integers(0, 92, (B, n_c)).astype(np.int32), rng.random(B, dtype=np.float32)
        d = lambda a: torch.tensor(a, device=dev)
        # --- the reference's loop body, fit.py:399-458 ---
        opt.inject_noise(d(us), d(ir), d(ua))
        opt.try_step()
        Eb = energy.view(-1, B)
        z = ((Eb - Eb.mean(-1).unsqueeze(-1)) / Eb.std(-1).unsqueeze(-1)).view(-1)
        opt.zero_grad()
        new_losses = g.calculate_energy(hand, obj, energy_fnc=energy_fnc, method="gendexgrasp", svd_gain=0.1)
        new_energy = _total(new_losses)
        new_energy.sum().backward()
        with torch.no_grad():
            accept, temp = opt.accept_step(energy, new_energy, None, z, 1.0)
            energy[accept] = new_energy.detach()[accept]
        # --- same iteration by the fused engine and by the oracle ---
        eng.step(d(us), d(ir), d(ua))
        out = orc.iterate(torch.tensor(us, dtype=torch.float64), torch.tensor(ir).long(), torch.tensor(ua, dtype=torch.float64))
        np.testing.assert_array_equal(accept.cpu().numpy(), out["accept"].numpy())
        np.testing.assert_array_equal(accept.cpu().numpy(), eng.accept.cpu().numpy().astype(bool))
        np.testing.assert_allclose(new_energy.detach().cpu().numpy(), out["new_energy"].numpy(), rtol=1e-4)
        np.testing.assert_allclose(energy.cpu().numpy(), eng.energy.cpu().numpy(), rtol=1e-4)
        np.testing.assert_allclose(hand.hand_pose.detach().cpu().numpy(), eng.pose.cpu().numpy(), rtol=1e-4, atol=1e-5)
        np.testing.assert_array_equal(hand.contact_point_indices.cpu().numpy(), eng.idx.cpu().numpy())
    eng.close()


def test_torchsdf_dropin(oracle_built):
    from graspqp_b200 import synth, torchsdf
    from oracle import sdf as osdf
    v, f = synth.make_mesh("torus", 2000, 1)
    fv = torchsdf.index_vertices_by_faces(torch.tensor(v), torch.tensor(f))
    pts = torch.tensor(np.random.default_rng(0).uniform(-0.15, 0.15, (5000, 3)).astype(np.float32), device="cuda",
                       requires_grad=True)
    d2, sg, nr, cl = torchsdf.compute_sdf(pts, fv.cuda())
    want = osdf.compute_sdf_port(pts.detach().cpu().numpy(), fv.numpy())
    for a, b in zip((d2, sg, nr, cl), want):
        np.testing.assert_array_equal(a.detach().cpu().numpy(), b)
    d2.sum().backward()  # tests/normal.py:36-39
    np.testing.assert_allclose(pts.grad.cpu().numpy(), (nr * 2 * d2.sqrt().unsqueeze(-1)).detach().cpu().numpy(), atol=5e-7)
