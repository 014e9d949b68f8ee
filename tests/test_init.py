"""(f1) initialize_convex_hull on the device (csrc/gqp_init.cu) vs its restatement (oracle/init.py) under injected noise;
the truncated normal against torch.nn.init.trunc_normal_ itself (the one piece of initializations.py:15-193 whose
package is present in this image)."""
import numpy as np
import pytest
import torch

from graspqp_b200 import engine, hand_pack, initializations, synth
from oracle import init as oinit


def test_trunc_normal_matches_torch():
    g = torch.Generator().manual_seed(5)
    t = torch.empty(4000)
    torch.nn.init.trunc_normal_(t, 0.3, 0.2, -0.1 - 1e-6, 0.9 + 1e-6, generator=g)
    u = torch.rand(4000, generator=torch.Generator().manual_seed(5)).numpy()
    np.testing.assert_allclose(oinit.trunc_normal(u, 0.3, 0.2, -0.1 - 1e-6, 0.9 + 1e-6), t.numpy(), atol=3e-6)


def test_hull_is_outward_and_closed():
    v, f = synth.make_mesh("torus", 2000, 3)
    tris = oinit.convex_hull(v)
    n = np.cross(tris[:, 1] - tris[:, 0], tris[:, 2] - tris[:, 0])
    assert ((n * (tris.mean(1) - v.mean(0))).sum(1) > 0).all()
    np.testing.assert_array_equal(tris, initializations.convex_hull(v))  # product copy of the same recipe


def check_device_init(lib, be, hand="allegro", B=16, n_c=12, pool_factor=20, seed=0):
    pack = hand_pack.load_pack(hand)
    hulls = [oinit.convex_hull(synth.make_mesh(k, 400, seed + i)[0]) for i, k in enumerate(("box", "hull"))]
    O, dof = len(hulls), len(pack["joints_lower"])
    rng = np.random.default_rng(seed)
    M = pool_factor * B
    u_pool = rng.random((O, M, 3), dtype=np.float32)
    u_pose = rng.random((O * B, 4 + dof), dtype=np.float32)
    idx_rand = rng.integers(0, len(pack["cand"]), (O * B, n_c)).astype(np.int32)
    want_pose, want_idx, picks = oinit.initialize_convex_hull(pack, hulls, B, n_c, u_pool, u_pose, idx_rand,
                                                              pool_factor=pool_factor)
    h = hand_pack.create_hand(lib, pack)
    di = initializations.DeviceInit(lib, be, pack, hulls, B, n_c, pool_factor=pool_factor)
    pose, idx, picked = be.empty((O * B, 9 + dof), np.float32), be.empty((O * B, n_c), np.int32), be.empty((O * B,), np.int32)
    di.run(h, pose, idx, be.from_host(u_pool), be.from_host(u_pose), be.from_host(idx_rand), picked=picked)
    np.testing.assert_array_equal(be.to_host(picked).reshape(O, B), np.stack(picks))  # FPS picks: bit-exact
    np.testing.assert_array_equal(be.to_host(idx), want_idx)
    got = be.to_host(pose)
    np.testing.assert_allclose(got[:, :3], want_pose[:, :3], atol=2e-7)   # translation
    np.testing.assert_allclose(got[:, 3:9], want_pose[:, 3:9], atol=2e-6)  # rotation columns
    np.testing.assert_allclose(got[:, 9:], want_pose[:, 9:], atol=3e-5)   # joints: two erfinv approximations, steep in the tails
    # sanity of the distribution: poses 5-10 cm (+1 cm inflation) off the hull, looking at it, joints inside their limits
    R0 = got[:, 3:6] / np.linalg.norm(got[:, 3:6], axis=1, keepdims=True)
    assert np.allclose(np.linalg.norm(got[:, 3:6], axis=1), 1, atol=1e-5) and np.abs((R0 * got[:, 6:9]).sum(1)).max() < 1e-5
    lo, hi = pack["joints_lower"], pack["joints_upper"]
    assert (got[:, 9:] >= lo - 2e-6).all() and (got[:, 9:] <= hi + 2e-6).all()
    # Philox path: deterministic in (seed, counter), different between calls
    p1, p2 = be.empty((O * B, 9 + dof), np.float32), be.empty((O * B, 9 + dof), np.float32)
    di.counter = 7
    di.run(h, p1, idx, seed=3)
    di.counter = 7
    di.run(h, p2, idx, seed=3)
    np.testing.assert_array_equal(be.to_host(p1), be.to_host(p2))
    di.run(h, p2, idx, seed=3)
    assert np.abs(be.to_host(p1) - be.to_host(p2)).max() > 1e-3
    assert np.isfinite(be.to_host(p2)).all()
    di.close()
    lib.gqp_hand_destroy(h)


@pytest.mark.parametrize("hand", ["allegro", "panda"])
def test_device_init_vs_oracle(emu, hand):
    check_device_init(emu, engine.NumpyBackend(), hand)


def test_engine_reset_step_with_device_init(emu):
    """FitEngine.reset_step: device init -> gqp_mala_reset -> evaluation -> forced accept of the masked rows"""
    pack = hand_pack.load_pack("allegro")
    v, f = synth.make_mesh("sphere", 200, 3)
    fv, sp = synth.face_vertices(v, f), synth.sample_surface(v, f, 48, seed=3)
    B, n_c = 4, 12
    pose, idx = synth.random_grasp_init(pack, [v], B, n_c, 3)
    eng = engine.FitEngine(emu, engine.NumpyBackend(), pack, [fv], sp[None], B, pose, idx.astype(np.int32))
    eng.enable_device_init([oinit.convex_hull(v)], pool_factor=8)
    eng.step()
    eng.reset_step(z_score_threshold=0.0)
    mask = eng.reset_mask.astype(bool)
    assert 0 < mask.sum() < B and int(eng.reset_count[0]) == mask.sum()
    np.testing.assert_array_equal(eng.pose[mask], eng.pose_init[mask])
    np.testing.assert_array_equal(eng.idx[mask], eng.idx_init[mask])
    assert (eng.step_count[mask] == 0).all() and (eng.step_count[~mask] == 2).all()
    assert (eng.ema[mask] == 0).all() and eng.accept[mask].all()
    eng.close()
