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
