"""FK / contacts / penetration / self-penetration kernels run by the host emulator vs the torch oracle (CPU only)."""
import ctypes
import time

import numpy as np
import pytest
import torch

from conftest import ptr
from graspqp_b200 import hand_pack, synth
from oracle.hand import HandOracle


@pytest.fixture(scope="module")
def pack():
    return hand_pack.load_pack("allegro")


@pytest.fixture(scope="module")
def hand(emu, pack):
    t = time.time()
    h = hand_pack.create_hand(emu, pack)
    print("hand_create %.2fs" % (time.time() - t))
    yield h
    emu.gqp_hand_destroy(h)


def _poses(pack, B, seed=0):
    v, f = synth.make_mesh("sphere", 200, 1)
    pose, idx = synth.random_grasp_init(pack, [v], B, 12, seed)
    return pose, idx.astype(np.int32)


def _oracle_world_tf(o):
    """world link transforms (B,L,12) from the oracle state"""
    out = []
    Rg, tg = o.global_rotation, o.global_translation
    for li in range(o.L):
        T = o.link_matrix(li)
        R = Rg @ T[:, :3, :3]
        t = (Rg @ T[:, :3, 3:]).squeeze(-1) + tg
        out.append(torch.cat([R, t.unsqueeze(-1)], -1).reshape(-1, 12))
    return torch.stack(out, 1)


def test_fk_forward_backward(emu, pack, hand):
    B = 5
    pose, idx = _poses(pack, B)
    L = len(pack["mesh_frame"])
    tf = np.zeros((B, L, 12), np.float32)
    assert emu.gqp_fk_forward(hand, ptr(pose), B, ptr(tf), None) == 0
    o = HandOracle(pack, torch.float64)
    p64 = torch.tensor(pose, dtype=torch.float64, requires_grad=True)
    o.set_parameters(p64, torch.tensor(idx).long())
    want = _oracle_world_tf(o)
    np.testing.assert_allclose(tf, want.detach().numpy(), atol=2e-6)
    # backward: random cotangent on the link transforms
    rng = np.random.default_rng(0)
    g = rng.normal(size=(B, L, 12)).astype(np.float32)
    extra = rng.normal(size=pose.shape).astype(np.float32)
    (want * torch.tensor(g, dtype=torch.float64)).sum().backward()
    gp = np.zeros_like(pose)
    assert emu.gqp_fk_backward(hand, ptr(pose), ptr(g), ptr(extra), B, ptr(gp), None) == 0
    np.testing.assert_allclose(gp - extra, p64.grad.numpy(), rtol=2e-4, atol=2e-5)


def test_contacts(emu, pack, hand):
    B, nc = 4, 12
    pose, idx = _poses(pack, B, 3)
    L = len(pack["mesh_frame"])
    tf = np.zeros((B, L, 12), np.float32)
    emu.gqp_fk_forward(hand, ptr(pose), B, ptr(tf), None)
    cp, cn = np.zeros((B, nc, 3), np.float32), np.zeros((B, nc, 3), np.float32)
    assert emu.gqp_contacts_forward(hand, ptr(tf), ptr(idx), B, nc, ptr(cp), ptr(cn), None) == 0
    o = HandOracle(pack, torch.float64)
    p64 = torch.tensor(pose, dtype=torch.float64, requires_grad=True)
    o.set_parameters(p64, torch.tensor(idx).long())
    np.testing.assert_allclose(cp, o.contact_points.detach().numpy(), atol=2e-6)
    np.testing.assert_allclose(cn, o.contact_normals.detach().numpy(), atol=2e-6)
    rng = np.random.default_rng(1)
    g1, g2 = rng.normal(size=cp.shape).astype(np.float32), rng.normal(size=cp.shape).astype(np.float32)
    ((o.contact_points * torch.tensor(g1, dtype=torch.float64)).sum()
     + (o.contact_normals * torch.tensor(g2, dtype=torch.float64)).sum()).backward()
    gtf = np.zeros((B, L, 12), np.float32)
    assert emu.gqp_contacts_backward(hand, ptr(idx), ptr(g1), ptr(g2), B, nc, ptr(gtf), None) == 0
    gp = np.zeros_like(pose)
    assert emu.gqp_fk_backward(hand, ptr(pose), ptr(gtf), None, B, ptr(gp), None) == 0
    np.testing.assert_allclose(gp, p64.grad.numpy(), rtol=2e-4, atol=2e-5)


def _object(seed=2, n=96):
    v, f = synth.make_mesh("sphere", 300, seed)
    return v, f, synth.sample_surface(v, f, n, seed=seed)


def test_hand_distance_and_pen(emu, pack, hand):
    """H4: full distances vs the oracle, culled E_pen == un-culled E_pen, gradient vs autograd."""
    B, n_pts = 3, 96
    v, f, sp = _object()
    pose, idx = synth.random_grasp_init(pack, [v], B, 12, 5)
    pose[:, :3] *= 0.55  # push the hand into the object so that some samples penetrate
    idx = idx.astype(np.int32)
    L = len(pack["mesh_frame"])
    tf = np.zeros((B, L, 12), np.float32)
    emu.gqp_fk_forward(hand, ptr(pose), B, ptr(tf), None)
    x = np.ascontiguousarray(sp[None], np.float32)  # (n_obj=1, n_pts, 3)
    dist = np.zeros((B, n_pts), np.float32)
    e_full, e_cull = np.zeros(B, np.float32), np.zeros(B, np.float32)
    scratch = np.zeros(int(emu.gqp_hand_distance_scratch_floats(1, B, n_pts, 0)), np.float32)
    g_full, g_cull = np.zeros((B, L, 12), np.float32), np.zeros((B, L, 12), np.float32)
    assert emu.gqp_hand_distance(hand, ptr(tf), ptr(x), 1, B, n_pts, 0, ptr(dist), ptr(e_full), ptr(scratch), 100.0,
                                 ptr(g_full), None) == 0, emu.gqp_last_error()
    assert emu.gqp_hand_distance(hand, ptr(tf), ptr(x), 1, B, n_pts, 0, None, ptr(e_cull), ptr(scratch), 100.0,
                                 ptr(g_cull), None) == 0
    o = HandOracle(pack, torch.float32)
    p32 = torch.tensor(pose, requires_grad=True)
    o.set_parameters(p32, torch.tensor(idx).long())
    xs = torch.tensor(sp).unsqueeze(0).repeat(B, 1, 1)
    want = o.cal_distance(xs)
    np.testing.assert_allclose(dist, want.detach().numpy(), atol=1e-5)
    e_want = torch.relu(want).sum(-1)
    assert (e_want > 0).any(), "test pose does not penetrate"
    np.testing.assert_allclose(e_full, e_want.detach().numpy(), rtol=1e-4, atol=1e-6)
    np.testing.assert_allclose(e_cull, e_full, rtol=1e-6, atol=1e-7)
    np.testing.assert_allclose(g_cull, g_full, rtol=1e-5, atol=1e-6)
    (100.0 * e_want.sum()).backward()
    gp = np.zeros_like(pose)
    assert emu.gqp_fk_backward(hand, ptr(pose), ptr(g_cull), None, B, ptr(gp), None) == 0
    np.testing.assert_allclose(gp, p32.grad.numpy(), rtol=2e-3, atol=2e-3)


def test_self_pen_joints(emu, pack, hand):
    B = 6
    pose, idx = _poses(pack, B, 7)
    pose[:, 9:] += np.random.default_rng(0).normal(size=pose[:, 9:].shape).astype(np.float32) * 0.6  # curl + exceed limits
    L = len(pack["mesh_frame"])
    tf = np.zeros((B, L, 12), np.float32)
    emu.gqp_fk_forward(hand, ptr(pose), B, ptr(tf), None)
    es, ej = np.zeros(B, np.float32), np.zeros(B, np.float32)
    gtf, gj = np.zeros((B, L, 12), np.float32), np.zeros_like(pose)
    assert emu.gqp_self_pen_joints(hand, ptr(tf), ptr(pose), B, ptr(es), ptr(ej), 10.0, 1.0, ptr(gtf), ptr(gj), None) == 0
    o = HandOracle(pack, torch.float64)
    p64 = torch.tensor(pose, dtype=torch.float64, requires_grad=True)
    o.set_parameters(p64, torch.tensor(idx.astype(np.int64)))
    want_s = o.self_penetration()
    th = o.hand_pose[:, 9:]
    want_j = ((th > o.joints_upper) * (th - o.joints_upper)).sum(-1) + ((th < o.joints_lower) * (o.joints_lower - th)).sum(-1)
    assert (want_s > 0).any() and (want_j > 0).any()
    np.testing.assert_allclose(es, want_s.detach().numpy(), rtol=1e-4, atol=1e-6)
    np.testing.assert_allclose(ej, want_j.detach().numpy(), rtol=1e-5, atol=1e-6)
    (10.0 * want_s.sum() + want_j.sum()).backward()
    gp = np.zeros_like(pose)
    assert emu.gqp_fk_backward(hand, ptr(pose), ptr(gtf), ptr(gj), B, ptr(gp), None) == 0
    np.testing.assert_allclose(gp, p64.grad.numpy(), rtol=1e-3, atol=1e-4)


@pytest.mark.parametrize("hand", ["allegro", "shadow_hand", "panda", "robotiq2", "robotiq3", "ability_hand"])
def test_culling_proof_outside_cells_have_positive_sign(emu, hand, oracle_built):
    """every SET cell of every link's outside grid, and the region outside the link's box: the reference's brute-force
    sign (TorchSDF arithmetic, oracle port) is +1 for the points drawn there (all set cells; the GPU test draws more
    points per cell)"""
    import cases
    from oracle import sdf as osdf
    sign = lambda pts, fv: osdf.compute_sdf_port(pts, fv)[1]
    n_links, n_cells, n_points = cases.check_culling_proof(emu, sign, hand, per_cell=2, n_far=5000)
    assert n_links >= 1 and n_points > 1000


@pytest.mark.parametrize("hand", ["allegro", "robotiq2", "ability_hand"])
def test_fused_warp_kernels_match_the_separate_ones(emu, hand):
    """gqp_fk_contacts_forward / _backward (a warp per candidate, lane = frame) == gqp_fk_forward + gqp_contacts_forward
    and gqp_contacts_backward + gqp_fk_backward (a thread per candidate)"""
    from graspqp_b200 import hand_pack, synth
    pack = hand_pack.load_pack(hand)
    h = hand_pack.create_hand(emu, pack)
    B, n_c = 5, 6
    v, _ = synth.make_mesh("sphere", 100, 1)
    pose, idx = synth.random_grasp_init(pack, [v], B, n_c, 4)
    pose, idx = np.ascontiguousarray(pose, np.float32), np.ascontiguousarray(idx, np.int32)
    L, D = len(pack["mesh_frame"]), pose.shape[1]
    tf1, tf2 = np.zeros((B, L, 12), np.float32), np.zeros((B, L, 12), np.float32)
    cp1, cn1, cp2, cn2 = [np.zeros((B, n_c, 3), np.float32) for _ in range(4)]
    assert emu.gqp_fk_forward(h, ptr(pose), B, ptr(tf1), None) == 0
    assert emu.gqp_contacts_forward(h, ptr(tf1), ptr(idx), B, n_c, ptr(cp1), ptr(cn1), None) == 0
    assert emu.gqp_fk_contacts_forward(h, ptr(pose), ptr(idx), B, n_c, ptr(tf2), ptr(cp2), ptr(cn2), None) == 0
    np.testing.assert_allclose(tf2, tf1, atol=1e-7)
    np.testing.assert_allclose(cp2, cp1, atol=1e-7)
    np.testing.assert_allclose(cn2, cn1, atol=1e-7)
    rng = np.random.default_rng(0)
    g_cp, g_cn = rng.normal(size=(B, n_c, 3)).astype(np.float32), rng.normal(size=(B, n_c, 3)).astype(np.float32)
    g_tf0 = rng.normal(size=(B, L, 12)).astype(np.float32)
    extra = rng.normal(size=(B, D)).astype(np.float32)
    g_tf = g_tf0.copy()
    gp1, gp2 = np.zeros((B, D), np.float32), np.zeros((B, D), np.float32)
    assert emu.gqp_contacts_backward(h, ptr(idx), ptr(g_cp), ptr(g_cn), B, n_c, ptr(g_tf), None) == 0
    assert emu.gqp_fk_backward(h, ptr(pose), ptr(g_tf), ptr(extra), B, ptr(gp1), None) == 0
    g_in = g_tf0.copy()
    assert emu.gqp_fk_contacts_backward(h, ptr(pose), ptr(idx), ptr(g_cp), ptr(g_cn), ptr(g_in), ptr(extra), B, n_c,
                                        ptr(gp2), None) == 0
    np.testing.assert_array_equal(g_in, g_tf0)  # read only
    np.testing.assert_allclose(gp2, gp1, rtol=2e-5, atol=2e-5 * np.abs(gp1).max())
    emu.gqp_hand_destroy(h)
